"""N > 1 host logic on CPU: ShardedVerifier's partition / all-gather / merge under gloo, world_size 2,
with the oracle's CPU restatement of the shard kernels as backend; result == unsharded oracle."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from duodecoding_b200.sharded import ShardedVerifier, shard_bounds


def _case(seed, B, gamma, V, sigma):
    g = torch.Generator().manual_seed(seed)
    tl = torch.randn(B, gamma + 1, V, generator=g) * 3
    dl = tl[:, :gamma] + sigma * torch.randn(B, gamma, V, generator=g)
    q = oracle.norm_logits(dl, 1.0, 0, 0.0)
    toks = torch.stack([oracle.sample(q[:, j], torch.empty(B, V).exponential_(1, generator=g))[:, 0]
                        for j in range(gamma)], 1)
    r = torch.rand(B, gamma, generator=g)
    noise = torch.empty(B, V).exponential_(1, generator=g)
    return dl, tl, toks, r, noise


def _worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle.sharded_port import OracleShardBackend
        ok = True
        for seed, sigma in [(1, 0.5), (2, 1.5)]:
            B, gamma, V = 5, 4, 1003
            dl, tl, toks, r, noise = _case(seed, B, gamma, V, sigma)
            sv = ShardedVerifier(V, backend=OracleShardBackend())
            got = sv.verify(sv.slice_of(dl), sv.slice_of(tl), toks, r, sv.slice_of(noise), 1.0)
            want = oracle.fused_verify(dl, tl, toks, r, noise, 1.0, 0, 0.0)
            ok &= torch.equal(got["accepted_len"], want["accepted_len"]) and torch.equal(got["token"], want["token"])
            ok &= torch.equal(got["accept_mask"], want["accept_mask"]) and sv.collectives == 2
            ok &= torch.allclose(got["selected_target_p"], want["selected_target_p"], rtol=1e-5, atol=1e-9)
        # top-p under sharding: stats + 4 levels + index cut + partials = 7 exchanges, same answer as the unsharded oracle
        for seed, sigma, T, top_p, d_top_p in [(3, 0.5, 1.0, 0.9, 0.9), (4, 1.5, 0.7, 0.5, 0.0)]:
            B, gamma, V = 5, 4, 1003
            dl, tl, toks, r, noise = _case(seed, B, gamma, V, sigma)
            sv = ShardedVerifier(V, backend=OracleShardBackend())
            got = sv.verify(sv.slice_of(dl), sv.slice_of(tl), toks, r, sv.slice_of(noise), T, top_p=top_p,
                            draft_top_p=d_top_p)
            want = oracle.fused_verify(dl, tl, toks, r, noise, T, 0, top_p, T, 0, d_top_p)
            ok &= torch.equal(got["accepted_len"], want["accepted_len"]) and torch.equal(got["token"], want["token"])
            ok &= torch.equal(got["accept_mask"], want["accept_mask"]) and sv.collectives == 7
            ok &= torch.allclose(got["selected_target_p"], want["selected_target_p"], rtol=5e-4, atol=1e-9)
            ok &= torch.allclose(got["selected_draft_p"], want["selected_draft_p"], rtol=5e-4, atol=1e-9)
        ret[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


def test_shard_bounds_cover_and_align():
    for V, world in [(151936, 8), (151936, 4), (32000, 2), (1003, 2), (7, 4)]:
        edges = [shard_bounds(V, g, world) for g in range(world)]
        assert edges[0][0] == 0 and edges[-1][1] == V
        assert all(a[1] == b[0] for a, b in zip(edges, edges[1:])) and all(lo % 8 == 0 for lo, _ in edges if lo < V)


def test_sharded_verify_world2_gloo():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(2, port, ret), nprocs=2, join=True)
    assert ret.get(0) is True and ret.get(1) is True


def _emulated_top_p(world, dl, tl, toks, r, noise, T, top_p, d_top_p):
    """ShardedVerifier's nucleus flow with the oracle backend, the all-gathers emulated by stacking."""
    from oracle.sharded_port import OracleShardBackend
    be = OracleShardBackend()
    B, gamma = toks.shape
    V = tl.shape[-1]
    parts = [(dl[..., a:b], tl[..., a:b], noise[..., a:b], a) for a, b in (shard_bounds(V, g, world) for g in range(world))]
    gstats = torch.stack([be.row_stats(d, t, toks, a, T, T) for d, t, _, a in parts])
    levels = be.nucleus_levels(parts[0][1], T, T)
    sels, quotas = [None] * world, [None] * world
    for level in range(levels):
        hists = []
        for g, (d, t, _, a) in enumerate(parts):
            sels[g], h = be.nucleus_level(d, t, gamma, a, T, T, top_p, d_top_p, gstats, sels[g], level)
            hists.append(h)
        gh = torch.stack(hists)
        for g in range(world):
            sels[g], quotas[g] = be.nucleus_reduce(gh, g, B, gamma, top_p, d_top_p, sels[g], level, levels, quotas[g])
    gcand = torch.stack([be.nucleus_tie(d, t, gamma, a, T, T, sels[g], quotas[g], levels)
                         for g, (d, t, _, a) in enumerate(parts)])
    cuts = be.nucleus_close(sels[0], gstats, gcand, B, gamma, levels)
    out, plan = be.accept(gstats, B, gamma, r, tokens=toks, cuts=cuts)
    partials = torch.stack([be.sample_partial(d, t, plan, nz, a, T, T, gamma, filtered=True) for d, t, nz, a in parts])
    out["token"] = be.finish(partials, B)
    return out, cuts


def test_sharded_top_p_protocol_on_cpu_with_ties():
    """The radix-select protocol (levels, reduction, index cut across ranks, close) in the oracle backend on rows
    full of equal values (bf16 grid), for 1, 2, 3 and 5 emulated ranks: the same filter for every partition, and the
    unsharded oracle's decisions."""
    for seed, T, top_p, d_top_p in [(11, 1.0, 0.9, 0.9), (12, 0.7, 0.6, 0.0), (13, 1.0, 0.999, 0.3)]:
        dl, tl, toks, r, noise = _case(seed, 4, 3, 1500, 1.0)
        dl, tl = (dl * 0.5).bfloat16().float(), (tl * 0.5).bfloat16().float()
        want = oracle.fused_verify(dl, tl, toks, r, noise, T, 0, top_p, T, 0, d_top_p)
        ref = None
        for world in (1, 2, 3, 5):
            got, cuts = _emulated_top_p(world, dl, tl, toks, r, noise, T, top_p, d_top_p)
            assert torch.equal(got["accepted_len"], want["accepted_len"]), (seed, world)
            assert torch.equal(got["token"], want["token"]) and torch.equal(got["accept_mask"], want["accept_mask"])
            assert torch.allclose(got["selected_target_p"], want["selected_target_p"], rtol=5e-4, atol=1e-9)
            c = cuts.clone()
            c[c[:, 3] == 2, 2] = 0  # rows without a filter: the float sum of partial sums depends on the partition
            ref = c if ref is None else ref
            assert torch.equal(c, ref), (seed, world)
