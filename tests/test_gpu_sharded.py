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


def _run_top_p(world, dl, tl, toks, r, noise, T, top_p, d_top_p, dev):
    """The nucleus flow of ShardedVerifier._step with the all-gathers emulated by stacking: every emulated rank
    keeps its own selection state, as the real ranks do."""
    from duodecoding_b200.sharded import _CudaBackend, shard_bounds
    be = _CudaBackend()
    B, gamma = toks.shape
    V = tl.shape[-1]
    toks_d, r_d = toks.to(dev), r.to(dev)
    sl = [shard_bounds(V, g, world) for g in range(world)]
    parts = [(dl[..., a:b].to(dev), tl[..., a:b].to(dev), noise[..., a:b].contiguous().to(dev), a) for a, b in sl]
    gstats = torch.stack([be.row_stats(d, t, toks_d, a, T, T) for d, t, _, a in parts])
    levels = be.nucleus_levels(parts[0][1], T, T)
    sels, quotas = [None] * world, [None] * world
    for level in range(levels):
        hists = []
        for g, (d, t, _, a) in enumerate(parts):
            sels[g], h = be.nucleus_level(d, t, gamma, a, T, T, top_p, d_top_p, gstats, sels[g], level)
            hists.append(h)
        ghist = torch.stack(hists)
        for g in range(world):
            sels[g], quotas[g] = be.nucleus_reduce(ghist, g, B, gamma, top_p, d_top_p, sels[g], level, levels, quotas[g])
    for g in range(1, world):
        assert torch.equal(sels[g], sels[0])  # every rank reaches the same selection state
    gcand = torch.stack([be.nucleus_tie(d, t, gamma, a, T, T, sels[g], quotas[g], levels)
                         for g, (d, t, _, a) in enumerate(parts)])
    cuts = be.nucleus_close(sels[0], gstats, gcand, B, gamma, levels)
    out, plan = be.accept(gstats, B, gamma, r_d, tokens=toks_d, cuts=cuts)
    partials = torch.stack([be.sample_partial(d, t, plan, nz, a, T, T, gamma, filtered=True) for d, t, nz, a in parts])
    out["token"] = be.finish(partials, B)
    torch.cuda.synchronize()
    return out, cuts, levels


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_shard_top_p_matches_unsharded_oracle_for_every_world_size(dtype):
    """Top-p under vocabulary sharding (shard_nucleus.cuh) against the unsharded oracle (src/utils.py:650-709,
    src/decoding_ops.py:109-220): lengths, tokens and masks exact, selected probabilities up to one boundary entry
    (cf. NUCLEUS_RTOL in test_gpu_parity.py); the filter itself is bit-identical for 1, 2, 3 and 8 ranks."""  # noqa
    dev = torch.device("cuda:0")
    for seed, sigma, T, top_p, d_top_p in [(1, 0.5, 1.0, 0.9, 0.9), (2, 1.5, 1.0, 0.5, 0.95), (3, 1.0, 0.7, 0.9, 0.9),
                                            (4, 1.0, 1.0, 0.9, 0.0), (5, 1.0, 1.0, 1.0, 0.3)]:
        dl, tl, toks, r, noise = _case(seed, 6, 4, 6000, sigma)
        if dtype == torch.bfloat16:  # coarse values: thousands of exact ties, the index cut decides
            dl, tl = (dl * 0.5).to(dtype), (tl * 0.5).to(dtype)
        want = oracle.fused_verify(dl.float(), tl.float(), toks, r, noise, T, 0, top_p, T, 0, d_top_p)
        ref_cuts = None
        for world in (1, 2, 3, 8):
            got, cuts, levels = _run_top_p(world, dl, tl, toks, r, noise, T, top_p, d_top_p, dev)
            assert levels == (4 if dtype == torch.bfloat16 and T == 1.0 else 8)
            what = (seed, world)
            assert torch.equal(got["accepted_len"].cpu(), want["accepted_len"]), what
            assert torch.equal(got["token"].cpu(), want["token"]), what
            assert torch.equal(got["accept_mask"].cpu(), want["accept_mask"]), what
            torch.testing.assert_close(got["selected_target_p"].cpu(), want["selected_target_p"], rtol=5e-4, atol=1e-9)
            torch.testing.assert_close(got["selected_draft_p"].cpu(), want["selected_draft_p"], rtol=5e-4, atol=1e-9)
            c = cuts.view(torch.int32).clone()  # [rows, 4]: x* bits, i*, kept-mass bits, state
            c[c[:, 3] == 2, 2] = 0  # (rows without a filter carry the float sum of the partial sums)
            if ref_cuts is None:
                ref_cuts = c
            assert torch.equal(c, ref_cuts), what  # integer masses: one answer for every partition


def test_sharded_verifier_world1_and_edge_cases():
    from duodecoding_b200.sharded import ShardedVerifier
    dev = torch.device("cuda:0")
    dl, tl, toks, r, noise = _case(7, 3, 2, 2048, 1.0)
    sv = ShardedVerifier(2048, rank=0, world=1)
    got = sv.verify(dl.to(dev), tl.to(dev), toks.to(dev), r.to(dev), noise.to(dev), 1.0)
    want = oracle.fused_verify(dl, tl, toks, r, noise, 1.0, 0, 0.0)
    assert torch.equal(got["token"].cpu(), want["token"]) and torch.equal(got["accepted_len"].cpu(), want["accepted_len"])
    for _ in range(2):  # top-p through the verifier (second call: the captured graph)
        got = sv.verify(dl.to(dev), tl.to(dev), toks.to(dev), r.to(dev), noise.to(dev), 1.0, top_p=0.9)
        want_p = oracle.fused_verify(dl, tl, toks, r, noise, 1.0, 0, 0.9)
        assert torch.equal(got["token"].cpu(), want_p["token"])
        assert torch.equal(got["accepted_len"].cpu(), want_p["accepted_len"])
        torch.testing.assert_close(got["selected_target_p"].cpu(), want_p["selected_target_p"], rtol=5e-4, atol=1e-9)
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


def test_shard_top_p_adversarial_rows_keep_the_oracles_support():
    """The filters (x*, i*) of the distributed select, decoded into kept sets, against the support of the oracle's
    top-p filter (float64 reference of src/utils.py:669-681) on rows built to stress it: all entries equal (the index
    cut runs across the ranks), two plateaus, -inf entries, one dominant entry, a tiny and a near-1 top_p."""
    dev = torch.device("cuda:0")
    B, gamma, V = 2, 3, 4104
    g = torch.Generator().manual_seed(12)
    tl = torch.randn(B, gamma + 1, V, generator=g) * 3
    tl[0, 0] = 0.25                                   # uniform: the first ceil(top_p * V) indices are kept
    tl[0, 1, : V // 2], tl[0, 1, V // 2:] = 1.5, -0.5  # two plateaus
    tl[0, 2, ::3] = float("-inf")                      # a third of the row masked out upstream
    tl[0, 3] = -5.0
    tl[0, 3, 1234] = 20.0                              # one dominant entry
    tl[1, 0] = (tl[1, 0] * 4).round() / 4              # coarse grid: many ties at the boundary
    dl = tl[:, :gamma] + 0.5 * torch.randn(B, gamma, V, generator=g)
    toks = torch.randint(0, V, (B, gamma), generator=g)
    r = torch.rand(B, gamma, generator=g)
    noise = torch.empty(B, V).exponential_(1, generator=g)
    idx = torch.arange(V)
    for top_p in (0.9, 0.02, 0.999):
        for world in (1, 3, 8):
            _, cuts, _ = _run_top_p(world, dl, tl, toks, r, noise, 1.0, top_p, top_p, dev)
            c = cuts.cpu().view(torch.int32).reshape(B, 2 * gamma + 1, 4)
            for b in range(B):
                for j in range(gamma + 1):
                    row = tl[b, j]
                    cut = c[b, 2 * j + 1 if j < gamma else 2 * gamma]
                    xs, istar = cut[:1].view(torch.float32).item(), int(cut[1])
                    got = (row > xs) | ((row == xs) & (idx <= istar))
                    want = torch.isfinite(oracle.filter_logits(row.double().unsqueeze(0), 0, top_p)[0])
                    diff = int((got != want).sum())
                    # the float64 oracle and the 40-bit fixed-point sums may disagree on entries at the boundary
                    # whose inclusion changes the kept mass by less than 1e-6
                    assert diff <= 2, (top_p, world, b, j, diff, int(got.sum()), int(want.sum()))
                    if diff:
                        p = torch.softmax(row.double(), 0)
                        assert float(p[got != want].sum()) < 1e-3 and abs(float(p[got].sum()) - float(p[want].sum())) < 1e-3
                    # equal values are kept lowest index first
                    ties = row == xs
                    if bool(ties.any()) and xs > float("-inf"):
                        kept_ties = idx[ties & got]
                        assert len(kept_ties) >= 1 and bool((idx[ties][: len(kept_ties)] == kept_ties).all())
