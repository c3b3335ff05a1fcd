"""GPU parity of the vocabulary-shard kernels (through the C ABI): `world` slices of one problem are
processed on ONE device, the two all-gathers are emulated by stacking, result == unsharded oracle."""
import pytest
import torch

import oracle
from tests.test_sharded_gloo import _case

pytestmark = pytest.mark.gpu


def _run(world, dl, tl, toks, r, noise, T, dev):
    from duodecoding_b200.sharded import _CudaBackend, shard_bounds
    be = _CudaBackend()
    B, gamma = toks.shape
    V = tl.shape[-1]
    toks_d, r_d = toks.to(dev), r.to(dev)
    sl = [shard_bounds(V, g, world) for g in range(world)]
    parts = [(dl[..., a:b].to(dev), tl[..., a:b].to(dev), noise[..., a:b].contiguous().to(dev), a) for a, b in sl]
    stats = torch.stack([be.row_stats(d, t, toks_d, a, T, T) for d, t, _, a in parts])
    out, plan = be.accept(stats, B, gamma, r_d)
    partials = torch.stack([be.sample_partial(d, t, plan, nz, a, T, T, gamma) for d, t, nz, a in parts])
    out["token"] = be.finish(partials, B)
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize("world", [1, 2, 8])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_shard_kernels_match_unsharded_oracle(world, dtype):
    dev = torch.device("cuda:0")
    for seed, sigma, T in [(1, 0.5, 1.0), (2, 1.5, 1.0), (3, 1.0, 0.7)]:
        dl, tl, toks, r, noise = _case(seed, 6, 4, 5003 if world < 8 else 6000, sigma)
        dl, tl = dl.to(dtype), tl.to(dtype)
        if T != 1.0:  # tokens must come from the tempered draft distribution to get a mix of accepts
            pass
        want = oracle.fused_verify(dl.float(), tl.float(), toks, r, noise, T, 0, 0.0)
        got = _run(world, dl, tl, toks, r, noise, T, dev)
        assert torch.equal(got["accepted_len"].cpu(), want["accepted_len"])
        assert torch.equal(got["token"].cpu(), want["token"])
        assert torch.equal(got["accept_mask"].cpu(), want["accept_mask"])
        assert torch.equal(got["all_accepted"].cpu(), want["accepted_len"] == 4)
        torch.testing.assert_close(got["selected_target_p"].cpu(), want["selected_target_p"], rtol=1e-5, atol=1e-9)
        torch.testing.assert_close(got["selected_draft_p"].cpu(), want["selected_draft_p"], rtol=1e-5, atol=1e-9)


def test_sharded_verifier_world1_and_edge_cases():
    from duodecoding_b200.sharded import ShardedVerifier
    dev = torch.device("cuda:0")
    dl, tl, toks, r, noise = _case(7, 3, 2, 2048, 1.0)
    sv = ShardedVerifier(2048, rank=0, world=1)
    got = sv.verify(dl.to(dev), tl.to(dev), toks.to(dev), r.to(dev), noise.to(dev), 1.0)
    want = oracle.fused_verify(dl, tl, toks, r, noise, 1.0, 0, 0.0)
    assert torch.equal(got["token"].cpu(), want["token"]) and torch.equal(got["accepted_len"].cpu(), want["accepted_len"])
    # identical models: every draft accepted, bonus token from the last target row
    got = sv.verify(tl[:, :2].to(dev), tl.to(dev), toks.to(dev), torch.full((3, 2), 0.999, device=dev), noise.to(dev))
    assert got["accepted_len"].tolist() == [2, 2, 2] and got["all_accepted"].all()
    with pytest.raises(ValueError):
        sv.verify(dl[..., :100].to(dev), tl[..., :100].to(dev), toks.to(dev), r.to(dev), noise[..., :100].to(dev))


@pytest.mark.parametrize("world", [1, 2, 8])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_sharded_topk_candidate_merge_matches_unsharded_rebuild(world, dtype):
    """transfer_top_k over vocabulary slices (merge of per-shard top-k candidates) == the unsharded operator
    (src/utils.py:751-784 on the oracle), ties included (bf16 grid: many equal probabilities at the k-th place)."""
    from duodecoding_b200.sharded import ShardedTopK, shard_bounds
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(3)
    V = 6000
    rows = 7
    p = torch.softmax(torch.randn(rows, V, generator=g) * 2.5, -1).to(dtype)
    p[3] = 0
    p[3, 17] = 1.0  # a one-hot row: k - 1 kept zeros are holes
    p[4] = torch.softmax(torch.randn(V, generator=g) * 2.5, -1).to(torch.bfloat16).to(dtype)  # heavy ties
    for k in (1, 100, 1024, 5999):
        want = oracle.rebuild_topk_uniform_probs(p.float(), k)
        sl = [shard_bounds(V, r, world) for r in range(world)]
        ranks = [ShardedTopK(V, rank=r, world=world) for r in range(world)]
        parts = [p[:, a:b].contiguous().to(dev) for a, b in sl]
        cands = [rk.local_candidates(x, k) for rk, x in zip(ranks, parts)]
        kmax = max(c[0].shape[1] for c in cands)
        pad = lambda t, v: torch.nn.functional.pad(t, (0, kmax - t.shape[1]), value=v)  # noqa: E731
        gv = torch.stack([pad(c[0], -1.0) for c in cands])
        gi = torch.stack([pad(c[1], 0x7fffffff) for c in cands])
        got = torch.cat([rk.apply(x, rk.merge(gv, gi, k)) for rk, x in zip(ranks, parts)], -1).cpu().float()
        # (atol: with k = V - 1 the single hole takes 1 - sum(kept), a cancellation of order 1 ulp of 1.0)
        torch.testing.assert_close(got, want, rtol=1e-5 if dtype == torch.float32 else 1e-2, atol=2e-7)
