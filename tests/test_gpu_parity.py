"""GPU parity: the sm_100a kernels (through the C ABI) against the CPU oracle and
the reference-generated golden fixtures.  Integer outputs (accepted lengths,
token ids, masks, support sets) must be bit-exact; probabilities agree within
1e-5 relative in fp32 (north_star tolerance), 1e-2 in bf16."""

import numpy as np
import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu

RTOL = 1e-5
ATOL = 1e-10


@pytest.fixture(scope="module")
def ops():
    import duodecoding_b200.ops as ops

    return ops


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda:0")


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a))


def _close(got, want, rtol=RTOL, atol=ATOL):
    torch.testing.assert_close(got.cpu().float(), want.float(), rtol=rtol, atol=atol)


# The nucleus boundary: the reference decides "cum > top_p" on a softmax whose fp32
# denominator carries ~1e-6..1e-5 of summation noise (vectorised fp32 sum over V
# entries), i.e. the boundary position is only defined up to that much cumulative
# mass.  The kernels decide it on EXACT bin masses.  On flat synthetic rows the
# entries next to the boundary weigh ~1e-5 themselves, so the two can disagree
# by an entry or two there; on peaked rows (what an LM emits) they cannot.
BOUNDARY_EPS = 6e-5


def _nucleus_parity(got, want, logits, T, k, p):
    """Supports equal except entries whose float64 cumulative position is within
    BOUNDARY_EPS of top_p; probabilities equal up to the normaliser that follows
    from the support; exact (1e-5) on rows whose support matches."""
    got, want = got.cpu().float().reshape(-1, got.shape[-1]), want.float().reshape(-1, want.shape[-1])
    x = (logits.float().reshape(-1, logits.shape[-1]) / T).double()
    flips = 0
    for row in range(got.shape[0]):
        g, w = got[row], want[row]
        diff = ((g == 0) != (w == 0)).nonzero().flatten()
        if diff.numel() == 0:
            torch.testing.assert_close(g, w, rtol=RTOL, atol=ATOL)
            continue
        flips += int(diff.numel())
        xr = x[row]
        if k > 0:
            kth = torch.topk(xr, min(k, xr.numel()))[0][-1]
            xr = xr.masked_fill(xr < kth, float("-inf"))
        pr = torch.softmax(xr, -1)
        srt, order = torch.sort(pr, descending=True, stable=True)
        cum = torch.cumsum(srt, 0)
        pos = torch.empty_like(order)
        pos[order] = torch.arange(order.numel())
        for i in diff.tolist():
            j = int(pos[i])
            before = float(cum[j - 1]) if j > 0 else 0.0
            assert abs(before - p) < BOUNDARY_EPS, (row, i, before, p)
        both = (g != 0) & (w != 0)
        ratio = (g[both].double() / w[both].double())
        assert float(ratio.max() / ratio.min() - 1.0) < 2e-5, "same distribution up to the normaliser"
    return flips


# ------------------------------------------------------------------ norm_logits
def test_norm_logits_golden(golden, ops, dev):
    for si in range(3):
        x = _t(golden[f"norm/{si}/logits"])
        for wi in range(8):
            T, k, p = golden[f"norm/{si}/{wi}/params"]
            want = _t(golden[f"norm/{si}/{wi}/probs"])
            got = ops.norm_logits(x.to(dev), float(T), int(k), float(p))
            assert got.shape == want.shape and got.dtype == want.dtype
            if T == 0 or p == 0:
                assert torch.equal(got.cpu() == 0, want == 0), f"support differs shape {si} warp {wi}"
                _close(got, want)
            else:
                _nucleus_parity(got, want, x, float(T), int(k), float(p))
    x = _t(golden["norm2d/logits"])
    got = ops.norm_logits(x.to(dev), 0.9, 20, 0.8)
    want = _t(golden["norm2d/probs"])
    _nucleus_parity(got, want, x, 0.9, 20, 0.8)


@pytest.mark.parametrize("team", [1, 2, 4, 8])
@pytest.mark.parametrize("warp", [(1.0, 0, 0.0), (1.0, 0, 0.9), (0.7, 100, 0.0), (1.3, 64, 0.8), (0.0, 0, 0.0)])
def test_norm_logits_teams(ops, dev, team, warp):
    g = torch.Generator().manual_seed(11)
    x = torch.randn(6, 5000, generator=g) * 3
    T, k, p = warp
    want = oracle.norm_logits(x, T, k, p)
    ops.set_team_size(team)
    try:
        got = ops.norm_logits(x.to(dev), T, k, p)
    finally:
        ops.set_team_size(0)
    if T == 0 or p == 0:
        assert torch.equal(got.cpu() == 0, want == 0)
        _close(got, want)
    else:
        _nucleus_parity(got, want, x, T, k, p)


def test_norm_logits_large_vocab(ops, dev):
    g = torch.Generator().manual_seed(12)
    for V in (32000, 151936):
        x = torch.randn(3, V, generator=g) * 3
        for T, k, p in [(1.0, 0, 0.0), (1.0, 0, 0.9), (0.8, 1024, 0.95)]:
            want = oracle.norm_logits(x, T, k, p)
            got = ops.norm_logits(x.to(dev), T, k, p)
            if p == 0:
                assert torch.equal(got.cpu() == 0, want == 0), (V, T, k, p)
                _close(got, want)
            else:
                _nucleus_parity(got, want, x, T, k, p)


def test_norm_logits_peaked_rows_are_exact(ops, dev):
    # LM-like rows: a few tokens carry the nucleus, the boundary margin dwarfs fp32 noise
    g = torch.Generator().manual_seed(15)
    for V in (32000, 151936):
        x = torch.randn(4, V, generator=g) * 2
        hot = torch.randint(0, V, (4, 40), generator=g)
        x.scatter_(1, hot, torch.rand(4, 40, generator=g) * 6 + 9)
        for T, k, p in [(1.0, 0, 0.9), (0.7, 0, 0.95), (1.0, 50, 0.8)]:
            want = oracle.norm_logits(x, T, k, p)
            got = ops.norm_logits(x.to(dev), T, k, p)
            assert torch.equal(got.cpu() == 0, want == 0), (V, T, k, p)
            _close(got, want)


def test_norm_logits_flat_rows_take_the_slow_path(ops, dev):
    # near-uniform rows put thousands of entries in the boundary band -> bisection path
    g = torch.Generator().manual_seed(13)
    x = torch.randn(4, 3000, generator=g) * 1e-3
    want = oracle.norm_logits(x, 1.0, 0, 0.5)
    got = ops.norm_logits(x.to(dev), 1.0, 0, 0.5)
    # boundary decided on exp values: allow the support to differ by entries of equal probability only
    diff = (got.cpu() == 0) != (want == 0)
    assert int(diff.sum()) <= 2 * x.shape[0]
    assert abs(float(got.sum()) - x.shape[0]) < 1e-3


def test_norm_logits_bf16(ops, dev):
    g = torch.Generator().manual_seed(14)
    x = (torch.randn(4, 32000, generator=g) * 3).bfloat16()
    for T, k, p in [(1.0, 0, 0.0), (1.0, 0, 0.9), (1.0, 50, 0.0)]:
        want = oracle.norm_logits(x.float(), T, k, p)
        got = ops.norm_logits(x.to(dev), T, k, p, out_dtype=torch.float32)
        if p == 0:
            assert torch.equal(got.cpu() == 0, want == 0), (T, k, p)
            _close(got, want)
        else:
            _nucleus_parity(got, want, x, T, k, p)
        got16 = ops.norm_logits(x.to(dev), T, k, p)
        assert got16.dtype == torch.bfloat16
        torch.testing.assert_close(got16.cpu().float(), got.cpu(), rtol=1e-2, atol=1e-6)


# ------------------------------------------------------------------ fused verify
def _check_verify(got, want, rtol=RTOL):
    assert torch.equal(got["accept_mask"].cpu(), want["accept_mask"])
    assert torch.equal(got["accepted_len"].cpu(), want["accepted_len"])
    assert torch.equal(got["token"].cpu(), want["token"])
    assert torch.equal(got["all_accepted"].cpu(), want["all_accepted"])
    _close(got["selected_target_p"], want["selected_target_p"], rtol=rtol)
    _close(got["selected_draft_p"], want["selected_draft_p"], rtol=rtol)


# flat synthetic rows under top-p: one boundary entry more or less moves every
# probability of the row by its weight (see BOUNDARY_EPS above)
NUCLEUS_RTOL = 5e-4


def test_verify_fused_golden(golden, ops, dev):
    for ci in range(7):
        pre = f"verify/{ci}/"
        T, k, p, _ = golden[pre + "params"]
        tl, dl = _t(golden[pre + "target_logits"]), _t(golden[pre + "draft_logits"])
        toks, r, noise = _t(golden[pre + "tokens"]), _t(golden[pre + "r"]), _t(golden[pre + "noise"])
        got = ops.verify_fused(dl.to(dev), tl.to(dev), toks.to(dev), r.to(dev), noise.to(dev), float(T), int(k),
                               float(p))
        assert torch.equal(got["accepted_len"].cpu(), _t(golden[pre + "accepted_len"])), ci
        assert torch.equal(got["token"].cpu(), _t(golden[pre + "token"])), ci
        assert torch.equal(got["accept_mask"].cpu(), _t(golden[pre + "accept_mask"])), ci
        _close(got["selected_target_p"], _t(golden[pre + "selected_target_p"]))
        _close(got["selected_draft_p"], _t(golden[pre + "selected_draft_p"]))


def _make_case(seed, B, gamma, V, sigma, T, k, p, dtype=torch.float32, Vd=None):
    g = torch.Generator().manual_seed(seed)
    Vd = V if Vd is None else Vd
    tl = (torch.randn(B, gamma + 1, V, generator=g) * 3).to(dtype)
    dl = (tl[:, :gamma, :Vd].float() + sigma * torch.randn(B, gamma, Vd, generator=g)).to(dtype)
    q = oracle.norm_logits(dl.float(), T, k, p)
    toks = torch.stack(
        [oracle.sample(q[:, j], torch.empty(B, Vd).exponential_(1, generator=g))[:, 0] for j in range(gamma)], 1
    )
    r = torch.rand(B, gamma, generator=g)
    noise = torch.empty(B, V).exponential_(1, generator=g)
    return dl, tl, toks, r, noise


@pytest.mark.parametrize("team", [-5, 0, 1, 2, 4, 8])  # -5: the fast path (row pairs), else the general team kernel
@pytest.mark.parametrize(
    "warp", [(1.0, 0, 0.0), (1.0, 0, 0.9), (0.8, 40, 0.0), (1.2, 200, 0.7), (0.0, 0, 0.0), (0.7, 0, 0.5)]
)
def test_verify_fused_vs_oracle(ops, dev, team, warp):
    T, k, p = warp
    dl, tl, toks, r, noise = _make_case(21, 6, 4, 6000, 0.7, T, k, p)
    want = oracle.fused_verify(dl, tl, toks, r, noise, T, k, p)
    ops.set_team_size(max(team, 0))
    ops.set_stream_path(-team if team < 0 else 0)
    try:
        got = ops.verify_fused(dl.to(dev), tl.to(dev), toks.to(dev), r.to(dev), noise.to(dev), T, k, p)
    finally:
        ops.set_team_size(0)
        ops.set_stream_path(True)
    _check_verify(got, want, NUCLEUS_RTOL if p > 0 else RTOL)


@pytest.mark.parametrize("stream", [5, 0])
@pytest.mark.parametrize("sigma", [0.3, 1.0, 2.0])
def test_verify_fused_vocab_32000(ops, dev, sigma, stream):
    dl, tl, toks, r, noise = _make_case(22, 8, 4, 32000, sigma, 1.0, 0, 0.9)
    want = oracle.fused_verify(dl, tl, toks, r, noise, 1.0, 0, 0.9)
    ops.set_stream_path(stream)
    try:
        got = ops.verify_fused(dl.to(dev), tl.to(dev), toks.to(dev), r.to(dev), noise.to(dev), 1.0, 0, 0.9)
    finally:
        ops.set_stream_path(True)
    _check_verify(got, want, NUCLEUS_RTOL)


@pytest.mark.parametrize("stream", [5, 0])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_verify_fused_peaked_rows_are_exact(ops, dev, stream, dtype):
    # LM-like rows: nucleus of a few tokens, boundary margins far above fp32 noise -> 1e-5 everywhere
    g = torch.Generator().manual_seed(28)
    B, gamma, V = 16, 4, 32000
    tl = torch.randn(B, gamma + 1, V, generator=g) * 2
    hot = torch.randint(0, V, (B, gamma + 1, 30), generator=g)
    tl.scatter_(2, hot, torch.rand(B, gamma + 1, 30, generator=g) * 6 + 9)
    dl = tl[:, :gamma] + 0.4 * torch.randn(B, gamma, V, generator=g)
    tl, dl = tl.to(dtype), dl.to(dtype)
    q = oracle.norm_logits(dl.float(), 1.0, 0, 0.9)
    toks = torch.stack(
        [oracle.sample(q[:, j], torch.empty(B, V).exponential_(1, generator=g))[:, 0] for j in range(gamma)], 1
    )
    r, noise = torch.rand(B, gamma, generator=g), torch.empty(B, V).exponential_(1, generator=g)
    want = oracle.fused_verify(dl.float(), tl.float(), toks, r, noise, 1.0, 0, 0.9)
    ops.set_stream_path(stream)
    try:
        got = ops.verify_fused(dl.to(dev), tl.to(dev), toks.to(dev), r.to(dev), noise.to(dev), 1.0, 0, 0.9)
    finally:
        ops.set_stream_path(True)
    _check_verify(got, want)
    assert 0 < int(got["accepted_len"].sum()) < B * gamma  # both outcomes occur


def test_verify_fused_qwen_vocab_bf16(ops, dev):
    dl, tl, toks, r, noise = _make_case(23, 3, 6, 151936, 0.5, 1.0, 0, 0.0, dtype=torch.bfloat16)
    want = oracle.fused_verify(dl.float(), tl.float(), toks, r, noise, 1.0, 0, 0.0)
    got = ops.verify_fused(dl.to(dev), tl.to(dev), toks.to(dev), r.to(dev), noise.to(dev), 1.0, 0, 0.0)
    _check_verify(got, want)


@pytest.mark.parametrize("stream", [5, 0])
@pytest.mark.parametrize(
    "B,gamma,V,T,p,dtype",
    [
        (1, 1, 1000, 1.0, 0.9, torch.float32),
        (1, 5, 4096, 0.5, 0.8, torch.float32),
        (2, 7, 8200, 1.0, 0.95, torch.float32),
        (5, 1, 32000, 2.0, 0.9, torch.float32),
        (7, 2, 1504, 1.3, 0.99, torch.float32),
        (9, 5, 1000, 1.0, 1.0, torch.float32),
        (2, 3, 1008, 1.0, 0.5, torch.bfloat16),
        (3, 2, 2056, 0.8, 0.3, torch.bfloat16),
        (4, 8, 520, 1.0, 0.05, torch.float32),
    ],
)
def test_verify_fused_shape_sweep(ops, dev, stream, B, gamma, V, T, p, dtype):
    # small and ragged shapes: one-row batches, long and short drafts, a last ring stage of a few
    # entries, top_p from "almost nothing" to 1.0, temperatures on both sides of 1
    dl, tl, toks, r, noise = _make_case(100 + B * 17 + gamma, B, gamma, V, 0.6, T, 0, p, dtype=dtype)
    want = oracle.fused_verify(dl.float(), tl.float(), toks, r, noise, T, 0, p)
    ops.set_stream_path(stream)
    try:
        got = ops.verify_fused(dl.to(dev), tl.to(dev), toks.to(dev), r.to(dev), noise.to(dev), T, 0, p)
    finally:
        ops.set_stream_path(True)
    _check_verify(got, want, NUCLEUS_RTOL if dtype == torch.float32 else 1e-2)


@pytest.mark.parametrize("stream", [5, 0])
def test_verify_fused_odd_batch_odd_gamma(ops, dev, stream):
    # B * (gamma + 2) odd: the workspace sections behind the token block start 8 bytes off a
    # 16-byte boundary unless the launcher rounds up; also an odd row count (one half of the last
    # row pair has no row)
    dl, tl, toks, r, noise = _make_case(35, 3, 3, 6000, 0.7, 1.0, 0, 0.9)
    want = oracle.fused_verify(dl, tl, toks, r, noise, 1.0, 0, 0.9)
    ops.set_stream_path(stream)
    try:
        got = ops.verify_fused(dl.to(dev), tl.to(dev), toks.to(dev), r.to(dev), noise.to(dev), 1.0, 0, 0.9)
    finally:
        ops.set_stream_path(True)
    _check_verify(got, want, NUCLEUS_RTOL)


@pytest.mark.parametrize("stream", [5])
@pytest.mark.parametrize("dtype,V,T,p", [(torch.bfloat16, 151936, 1.0, 0.9), (torch.float32, 128256, 0.7, 0.95)])
def test_verify_fused_large_vocab_top_p(ops, dev, stream, dtype, V, T, p):
    # 75 / 63 ring stages per row, a ragged last stage and (T != 1) the IEEE division on the fast paths
    dl, tl, toks, r, noise = _make_case(31, 3, 3, V, 0.5, T, 0, p, dtype=dtype)
    want = oracle.fused_verify(dl.float(), tl.float(), toks, r, noise, T, 0, p)
    ops.set_stream_path(stream)
    try:
        got = ops.verify_fused(dl.to(dev), tl.to(dev), toks.to(dev), r.to(dev), noise.to(dev), T, 0, p)
    finally:
        ops.set_stream_path(True)
    _check_verify(got, want, NUCLEUS_RTOL if dtype == torch.float32 else 1e-2)


@pytest.mark.parametrize("stream", [5, 0])
def test_verify_fused_flat_rows_overflow_the_lists(ops, dev, stream):
    # nearly flat rows: the boundary bucket holds most of the vocabulary, far more than the
    # shared-memory lists take, so every row goes through the general search
    g = torch.Generator().manual_seed(33)
    B, gamma, V = 4, 3, 32000
    tl = torch.randn(B, gamma + 1, V, generator=g) * 0.05
    dl = tl[:, :gamma] + 0.02 * torch.randn(B, gamma, V, generator=g)
    q = oracle.norm_logits(dl, 1.0, 0, 0.9)
    toks = torch.stack(
        [oracle.sample(q[:, j], torch.empty(B, V).exponential_(1, generator=g))[:, 0] for j in range(gamma)], 1
    )
    r, noise = torch.rand(B, gamma, generator=g), torch.empty(B, V).exponential_(1, generator=g)
    want = oracle.fused_verify(dl, tl, toks, r, noise, 1.0, 0, 0.9)
    ops.set_stream_path(stream)
    try:
        got = ops.verify_fused(dl.to(dev), tl.to(dev), toks.to(dev), r.to(dev), noise.to(dev), 1.0, 0, 0.9)
    finally:
        ops.set_stream_path(True)
    _check_verify(got, want, NUCLEUS_RTOL)


def test_verify_fused_unequal_vocab(ops, dev):
    # draft vocabulary smaller than the target's: gathers and the residual use the common prefix
    dl, tl, toks, r, noise = _make_case(24, 4, 3, 1200, 0.8, 1.0, 0, 0.0, Vd=1000)
    want = oracle.fused_verify(dl, tl, toks, r, noise, 1.0, 0, 0.0)
    got = ops.verify_fused(dl.to(dev), tl.to(dev), toks.to(dev), r.to(dev), noise.to(dev), 1.0, 0, 0.0)
    _check_verify(got, want)


def test_verify_fused_strided_history_views(ops, dev):
    # rows taken out of a longer [B, L, V] history buffer, as KVCacheModel holds them
    g = torch.Generator().manual_seed(25)
    B, gamma, V, L = 3, 4, 2000, 12
    hist_t = torch.randn(B, L, V, generator=g) * 3
    hist_d = hist_t + 0.5 * torch.randn(B, L, V, generator=g)
    tl, dl = hist_t[:, 5 : 5 + gamma + 1], hist_d[:, 5 : 5 + gamma]
    q = oracle.norm_logits(dl, 1.0, 0, 0.0)
    toks = torch.stack(
        [oracle.sample(q[:, j], torch.empty(B, V).exponential_(1, generator=g))[:, 0] for j in range(gamma)], 1
    )
    r, noise = torch.rand(B, gamma, generator=g), torch.empty(B, V).exponential_(1, generator=g)
    want = oracle.fused_verify(dl, tl, toks, r, noise)
    ht, hd = hist_t.to(dev), hist_d.to(dev)
    got = ops.verify_fused(hd[:, 5 : 5 + gamma], ht[:, 5 : 5 + gamma + 1], toks.to(dev), r.to(dev), noise.to(dev))
    _check_verify(got, want)


def test_verify_fused_edge_cases(ops, dev):
    # r == 0 accepts everything (bonus path); r > 1 rejects at position 0
    dl, tl, toks, r, noise = _make_case(26, 4, 4, 3000, 0.5, 1.0, 0, 0.0)
    for rv in (0.0, 2.0):
        rr = torch.full_like(r, rv)
        want = oracle.fused_verify(dl, tl, toks, rr, noise)
        got = ops.verify_fused(dl.to(dev), tl.to(dev), toks.to(dev), rr.to(dev), noise.to(dev))
        _check_verify(got, want)
    # identical draft and target: p/q == 1 accepts for every r <= 1
    want = oracle.fused_verify(tl[:, :4], tl, toks, r, noise)
    got = ops.verify_fused(tl[:, :4].to(dev), tl.to(dev), toks.to(dev), r.to(dev), noise.to(dev))
    _check_verify(got, want)
    assert bool(got["all_accepted"].all())
    # gamma == 0: only the bonus token
    got = ops.verify_fused(dl[:, :0].to(dev), tl[:, :1].to(dev), toks[:, :0].to(dev), r[:, :0].to(dev), noise.to(dev))
    want = oracle.fused_verify(dl[:, :0], tl[:, :1], toks[:, :0], r[:, :0], noise)
    assert torch.equal(got["token"].cpu(), want["token"]) and int(got["accepted_len"].sum()) == 0


def test_verify_fused_repeatable_and_workspace_self_cleaning(ops, dev):
    dl, tl, toks, r, noise = _make_case(27, 5, 4, 4096, 1.0, 1.0, 0, 0.9)
    args = [t.to(dev) for t in (dl, tl, toks, r, noise)]
    first = ops.verify_fused(*args, 1.0, 0, 0.9)
    first = {k: v.clone() for k, v in first.items()}
    for it in range(10):
        ops.set_stream_path([0, 5][it % 2])  # the device paths share (and must restore) the workspace
        again = ops.verify_fused(*args, 1.0, 0, 0.9)
        ops.set_stream_path(True)
        for k in first:
            if first[k].is_floating_point():
                # the general team kernel uses expf, the row-stream kernel 2^(x*log2e - M0) with an exponent
                # offset M0 predicted from earlier rows: probabilities repeat to ~1e-6, not bitwise
                torch.testing.assert_close(first[k], again[k], rtol=RTOL, atol=1e-9)
            else:
                assert torch.equal(first[k], again[k]), k


# ------------------------------------------------------------------ probability-space ops
def test_sample_and_max_fn_golden(golden, ops, dev):
    p, e = _t(golden["sample/probs"]), _t(golden["sample/noise"])
    got = ops.sample_rows(p.to(dev), e.to(dev))
    assert torch.equal(got.cpu(), _t(golden["sample/idx"]))
    x = _t(golden["maxfn/x"])
    got = ops.max_fn(x.to(dev))
    want = _t(golden["maxfn/y"])
    assert torch.equal(got.cpu() == 0, want == 0)
    _close(got, want)


def test_residual_sample_vs_oracle(ops, dev):
    g = torch.Generator().manual_seed(31)
    P = torch.softmax(torch.randn(16, 5000, generator=g) * 3, -1)
    Q = torch.softmax(torch.randn(16, 5000, generator=g) * 3, -1)
    Q[3] = P[3]  # zero residual mass -> argmax fallback (index 0)
    e = torch.empty(16, 5000).exponential_(1, generator=g)
    want = oracle.sample_reject_token(P, Q, e)
    got = ops.residual_sample(P.to(dev), Q.to(dev), e.to(dev))
    assert torch.equal(got.cpu(), want)
    want = oracle.sample_accept_token(P, e)
    got = ops.sample_rows(P.to(dev), e.to(dev))
    assert torch.equal(got.cpu(), want)


def test_accept_probs_vs_oracle(ops, dev):
    g = torch.Generator().manual_seed(32)
    B, gamma, V = 5, 6, 999
    P = torch.softmax(torch.randn(B, gamma, V, generator=g) * 2, -1)
    Q = torch.softmax(torch.randn(B, gamma, V, generator=g) * 2, -1)
    toks = torch.randint(0, V, (B, gamma), generator=g)
    Q[0, 1, toks[0, 1]] = 0.0  # q == 0, p > 0 -> inf -> accept
    P[1, 0, toks[1, 0]] = 0.0
    Q[1, 0, toks[1, 0]] = 0.0  # 0/0 -> nan -> reject
    r = torch.rand(B, gamma, generator=g)
    want = oracle.accept_rows(Q, P, toks, r)
    got = ops.accept_probs(Q.to(dev), P.to(dev), toks.to(dev), r.to(dev))
    assert torch.equal(got["accept_mask"].cpu(), want["accept_mask"])
    assert torch.equal(got["accepted_len"].cpu(), want["accepted_len"])
    assert torch.equal(got["selected_draft_p"].cpu(), want["selected_draft_p"])
    assert torch.equal(got["selected_target_p"].cpu(), want["selected_target_p"])


# ------------------------------------------------------------------ transfer_top_k compression
def test_topk_rebuild_golden(golden, ops, dev):
    pr = _t(golden["rebuild/probs"])
    for k in (1, 7, 64, 999):
        got = ops.topk_rebuild(pr.to(dev), k, 0)
        # the uniform fill is (1 - sum(top-k)) / n_zero: for k close to V that is a cancellation
        # of two numbers near 1, defined only up to the fp32 summation order (a few 2^-24)
        _close(got, _t(golden[f"rebuild/k{k}"]), atol=ATOL if k < 999 else 2.4e-7)
    prn = _t(golden["rebuild_sparse/probs"])
    _close(ops.topk_rebuild(prn.to(dev), 512, 0), _t(golden["rebuild_sparse/k512"]))
    v1 = _t(golden["comm/probs1d"])
    for k in (1, 32):
        got = ops.topk_rebuild(v1.to(dev)[None], k, 1)[0]
        assert torch.equal(got.cpu(), _t(golden[f"comm/topk{k}"]))
    comp = _t(golden["comm/topk32"])
    _close(ops.topk_rebuild(comp.to(dev)[None], 0, 3)[0], _t(golden["comm/rebuild_full32"]))
    p3 = _t(golden["comm/probs3d"])
    _close(ops.topk_rebuild(p3.to(dev), 16, 2), _t(golden["comm/compress_rebuild16"]))


@pytest.mark.parametrize("team", [0, 1, 2, 8])  # 0: the one-block-per-row select kernel (default)
def test_topk_rebuild_ties_lowest_index_first(ops, dev, team):
    g = torch.Generator().manual_seed(33)
    # bf16-grid probabilities: many exact ties at the k-th value
    p = torch.softmax((torch.randn(5, 32000, generator=g) * 3).bfloat16().float(), -1)
    ops.set_team_size(team)
    try:
        for k in (100, 1024):
            want = oracle.rebuild_topk_probs(p, k)
            got = ops.topk_rebuild(p.to(dev), k, 0)
            assert torch.equal((got.cpu() - want).abs() > 1e-6, torch.zeros_like(want, dtype=torch.bool))
            _close(got, want)
            want1 = torch.stack([oracle.apply_top_k_compression(row, k) for row in p])
            got1 = ops.topk_rebuild(p.to(dev), k, 1)
            assert torch.equal(got1.cpu(), want1)
    finally:
        ops.set_team_size(0)


def test_topk_select_kernel_edges(ops, dev):
    """The one-block-per-row select kernel (select_kernels.cuh): ragged V (scalar tail, unaligned rows), k at both
    ends, rows with thousands of equal values, bf16 rows, the comm modes - against the oracle restatements of
    src/utils.py:751-784 and src/communication.py:288-395."""
    g = torch.Generator().manual_seed(91)
    for V in (5, 33, 1000, 4099, 32001):
        p = torch.softmax(torch.randn(6, V, generator=g) * 3, -1)
        p[1] = 1.0 / V  # every entry equal: the index cut decides everything
        p[2, : V // 2] = p[2, 0]  # a long run of equal values
        p[2] /= p[2].sum()
        p[3, V // 3:] = 0.0  # exact zeros below the k-th value
        for k in sorted({1, 2, V // 2, V - 1}):
            if k <= 0 or k >= V:
                continue
            want1 = torch.stack([oracle.apply_top_k_compression(row, k) for row in p])
            got1 = ops.topk_rebuild(p.to(dev), k, 1)
            assert torch.equal(got1.cpu(), want1), (V, k)
            _close(ops.topk_rebuild(p.to(dev), k, 0), oracle.rebuild_topk_probs(p, k), atol=3e-7)
            want2 = oracle.compress_rebuild_probs(p[None], k)[0]
            _close(ops.topk_rebuild(p.to(dev), k, 2), want2, atol=3e-7)
        comp = torch.stack([oracle.apply_top_k_compression(row, max(1, V // 4)) for row in p])
        _close(ops.topk_rebuild(comp.to(dev), 0, 3), oracle.rebuild_full_probs(comp), atol=3e-7)
    # rows of a strided [B, S, V] view (row stride != V) and bf16 storage
    p3 = torch.softmax(torch.randn(3, 5, 4096, generator=g) * 2, -1)
    view = p3.to(dev)[:, 1:4]
    got = ops.topk_rebuild(view, 64, 1)
    want = torch.stack([oracle.apply_top_k_compression(row, 64) for row in p3[:, 1:4].reshape(-1, 4096)]).view(3, 3, 4096)
    assert torch.equal(got.cpu(), want)
    pb = torch.softmax(torch.randn(4, 32000, generator=g) * 3, -1).bfloat16()
    gotb = ops.topk_rebuild(pb.to(dev), 1024, 1)
    wantb = torch.stack([oracle.apply_top_k_compression(row, 1024) for row in pb.float()]).bfloat16()
    assert torch.equal(gotb.cpu(), wantb)
    _close(ops.topk_rebuild(pb.to(dev), 1024, 0), oracle.rebuild_topk_probs(pb.float(), 1024).bfloat16(), rtol=1e-2, atol=1e-8)


def test_host_verifier_matches_device_call(ops, dev):
    """duodecoding_b200.host.HostVerifier: the pinned-arena step (one H2D copy, one D2H copy) and the per-tensor
    ``verify`` give exactly what ops.verify_fused gives on device tensors."""
    from duodecoding_b200.host import HostVerifier

    B, gamma, V = 8, 4, 32000
    draft, target, tokens, r, noise = _make_case(77, B, gamma, V, 1.0, 1.0, 0, 0.9)
    want = {k: v.cpu() for k, v in ops.verify_fused(*(t.to(dev) for t in (draft, target, tokens, r, noise)), 1.0, 0, 0.9).items()}
    for nsplit in (1, 3):
        hv = HostVerifier(B, gamma, V, torch.float32, dev, nsplit=nsplit)
        hs = hv.host_set()
        hs["draft"].copy_(draft[:, :gamma]), hs["target"].copy_(target), hs["tokens"].copy_(tokens)
        hs["r"].copy_(r), hs["noise"].copy_(noise)
        for _ in range(2):  # the second step reuses the arenas
            got = hv.verify_set(hs, True, 1.0, 0, 0.9)
            for k in ("accepted_len", "token", "accept_mask", "all_accepted"):
                assert torch.equal(got[k], want[k]), (nsplit, k)
            torch.testing.assert_close(got["selected_target_p"], want["selected_target_p"], rtol=NUCLEUS_RTOL, atol=1e-9)
        assert hv.h2d_bytes == (B * (2 * gamma + 1) * V + B * V) * 4 + B * gamma * 12
        got = hv.verify(draft.pin_memory(), target.pin_memory(), tokens, r, noise, 1.0, 0, 0.9)
        for k in ("accepted_len", "token", "accept_mask", "all_accepted"):
            assert torch.equal(got[k], want[k]), (nsplit, k)
        no_noise = hv.verify_set(hs, False, 1.0, 0, 0.9)  # device-drawn Exp(1): the accept side does not depend on it
        assert torch.equal(no_noise["accepted_len"], want["accepted_len"])


def test_kv_truncate(ops, dev):
    lens = torch.tensor([10, 3, 7, 0], dtype=torch.int32, device=dev)
    ops.kv_truncate(lens, 5)
    assert lens.tolist() == [5, 3, 5, 0]
    ops.kv_truncate(lens, torch.tensor([9, 1, 2, 4]))
    assert lens.tolist() == [5, 1, 2, 0]


def test_cpu_tensors_are_rejected(ops):
    with pytest.raises(RuntimeError):
        ops.norm_logits(torch.randn(2, 10), 1.0)


def test_verify_fused_on_two_streams_does_not_share_a_workspace(ops, dev):
    """Calls with the same (B, gamma) issued on different CUDA streams may overlap on the device: each stream
    gets its own workspace (records, per-sequence cells, hints), so the results equal the serial ones."""
    cases = [_make_case(51 + i, 16, 4, 32000, 1.0, 1.0, 0, 0.9) for i in range(2)]
    args = [[t.to(dev) for t in c] for c in cases]
    serial = [{k: v.clone() for k, v in ops.verify_fused(*a, 1.0, 0, 0.9).items()} for a in args]
    torch.cuda.synchronize()
    streams = [torch.cuda.Stream(device=dev) for _ in range(2)]
    outs = [[], []]
    for it in range(20):
        for s in range(2):
            with torch.cuda.stream(streams[s]):
                outs[s].append({k: v.clone() for k, v in ops.verify_fused(*args[s], 1.0, 0, 0.9).items()})
    torch.cuda.synchronize()
    keys = {k[:2] for k in ops._WORKSPACES}
    assert len({k for k in ops._WORKSPACES if k[2:] == (16, 4)}) >= 3, keys  # default stream + the two streams
    for s in range(2):
        for got in outs[s]:
            for k in ("accepted_len", "token", "accept_mask"):
                assert torch.equal(got[k], serial[s][k]), (s, k)
            # (each workspace carries its own exponent-offset hint: a boundary entry may flip, cf. NUCLEUS_RTOL)
            torch.testing.assert_close(got["selected_target_p"], serial[s]["selected_target_p"], rtol=NUCLEUS_RTOL, atol=1e-9)


def test_bf16_reference_goldens_decisions_agree_outside_bf16_rounding(ops, dev):
    """tests/golden/bf16_v1.npz holds what the UNMODIFIED reference computes IN BFLOAT16 (its models' dtype,
    src/engine.py:261): bf16 logits / T, bf16 softmax, bf16 p/q, bf16 comparison.  The kernels read the same bf16
    logits and compute in fp32 (canonicalisations): selected probabilities agree to bf16 precision, and
    every accept decision agrees unless |p/q - r| lies inside bf16 rounding of the ratio."""
    from tests.bf16_check import check_against_bf16_goldens

    def run(dl, tl, toks, r, T, top_k, top_p):
        noise = torch.ones(toks.shape[0], tl.shape[-1], device=dev)
        got = ops.verify_fused(dl.to(dev), tl.to(dev), toks.to(dev), r.to(dev), noise, T, top_k, top_p)
        return {k: v.cpu() for k, v in got.items()}

    check_against_bf16_goldens(run)
