"""Pin the oracle: it must reproduce, bit for bit, what the unmodified reference
produced on the seeded inputs stored by oracle/gen_golden.py."""
import numpy as np
import torch

import oracle as O


def T(a):
    return torch.from_numpy(np.asarray(a))


def same(a: torch.Tensor, b) -> bool:
    b = T(b)
    return a.shape == b.shape and a.dtype == b.dtype and torch.equal(torch.nan_to_num(a), torch.nan_to_num(b))


def test_norm_logits_matches_reference_bitwise(golden):
    n = 0
    for si in range(3):
        x = T(golden[f"norm/{si}/logits"])
        for wi in range(8):
            t, k, p = golden[f"norm/{si}/{wi}/params"]
            got = O.norm_logits(x, float(t), int(k), float(p))
            assert same(got, golden[f"norm/{si}/{wi}/probs"]), (si, wi)
            n += 1
    assert n == 24
    assert same(O.norm_logits(T(golden["norm2d/logits"]), 0.9, 20, 0.8), golden["norm2d/probs"])


def test_sample_and_max_fn(golden):
    idx = O.sample(T(golden["sample/probs"]), T(golden["sample/noise"]))
    assert same(idx, golden["sample/idx"])
    assert int(idx[1]) == 0 and int(idx[4]) == 0  # zero-mass rows fall back to index 0
    assert same(O.max_fn(T(golden["maxfn/x"])), golden["maxfn/y"])


def test_rebuild_topk(golden):
    pr = T(golden["rebuild/probs"])
    for k in (1, 7, 64, 999, 1000, 0):
        assert same(O.rebuild_topk_uniform_probs(pr, k), golden[f"rebuild/k{k}"]), k
    assert same(O.rebuild_topk_uniform_probs(T(golden["rebuild_sparse/probs"]), 512), golden["rebuild_sparse/k512"])


def test_communication_helpers(golden):
    v1 = T(golden["comm/probs1d"])
    for k in (1, 32, 600):
        assert same(O.apply_top_k_compression(v1, k), golden[f"comm/topk{k}"])
    assert same(O.rebuild_full_probs(O.apply_top_k_compression(v1, 32)), golden["comm/rebuild_full32"])
    p3 = T(golden["comm/probs3d"])
    assert same(O.compress_rebuild_probs(p3, 16), golden["comm/compress_rebuild16"])
    kth = torch.topk(p3, 16, dim=-1)[0][..., -1:]
    assert same(O.rebuild_full_probs(torch.where(p3 >= kth, p3, torch.zeros_like(p3))), golden["comm/rebuild_full3d"])


def test_fused_verify_rows_match_reference(golden):
    for ci in range(7):
        pre = f"verify/{ci}/"
        t, k, p, prefix_len = golden[pre + "params"]
        out = O.fused_verify(
            T(golden[pre + "draft_logits"]),
            T(golden[pre + "target_logits"]),
            T(golden[pre + "tokens"]),
            T(golden[pre + "r"]),
            T(golden[pre + "noise"]),
            temperature=float(t),
            top_k=int(k),
            top_p=float(p),
        )
        assert same(out["accepted_len"], golden[pre + "accepted_len"]), ci
        assert same(out["token"], golden[pre + "token"]), ci
        assert same(out["accept_mask"], golden[pre + "accept_mask"]), ci
        assert same(out["selected_target_p"], golden[pre + "selected_target_p"]), ci
        assert same(out["selected_draft_p"], golden[pre + "selected_draft_p"]), ci
        n = int(prefix_len) + out["accepted_len"] - 1
        assert same(n, golden[pre + "n"]), ci


def test_payload_truncation(golden):
    ph, x = T(golden["payload/ph"]), T(golden["payload/x"])
    for name in ("full", "rows", "toks", "none", "g0"):
        pl, gm = golden[f"payload/{name}/args"]
        tk, pq = O.collect_verification_payload(ph, x, int(pl), int(gm))
        assert same(tk, golden[f"payload/{name}/tokens"]), name
        assert same(pq, golden[f"payload/{name}/probs"]), name


def test_finalize_and_resolve_positions(golden):
    P, Q = T(golden["finalize/P"]), T(golden["finalize/Q"])
    x, noise = T(golden["finalize/x"]), T(golden["finalize/noise"])
    for name in ("reject", "accept"):
        prefix_len, gam, n = (int(v) for v in golden[f"finalize/{name}/args"])
        plan = O.build_rollback_plan(prefix_len, gam, n)
        if plan.all_accepted:
            t = O.sample_accept_token(P[:, -1], noise)
            ends = [plan.draft_end_pos, plan.target_end_pos_accept]
        else:
            t = O.sample_reject_token(P[:, n], Q[:, n], noise)
            ends = [plan.draft_end_pos, plan.target_end_pos_reject]
        y = torch.cat((x[:, : n + 1], t), dim=1)
        assert same(y, golden[f"finalize/{name}/out"])
        assert ends == golden[f"finalize/{name}/rollbacks"].tolist()
    prefix_len, gam = 8, 4
    for name in ("mixed", "all"):
        r = T(golden[f"resolve/{name}/r"])
        q = Q[:, prefix_len - 1 : prefix_len + gam - 1]
        p = P[:, prefix_len - 1 : prefix_len + gam]
        toks = x[:, prefix_len : prefix_len + gam]
        out = O.fused_verify(q, p, toks, r, noise, inputs_are_probs=True)
        # resolve_stage_verification samples the bonus from prob_history[:, -1]
        a = int(out["accepted_len"][0])
        n = prefix_len + a - 1
        plan = O.build_rollback_plan(prefix_len, gam, n)
        tok = int(O.sample_accept_token(P[:, -1], noise)) if plan.all_accepted else int(out["token"][0])
        want = golden[f"resolve/{name}/result"].tolist()
        assert [a, n, tok, int(plan.all_accepted)] == want
        t_end = plan.target_end_pos_accept if plan.all_accepted else plan.target_end_pos_reject
        assert [plan.draft_end_pos, t_end] == golden[f"resolve/{name}/rollbacks"].tolist()


def test_cuhlm_oracle_is_pinned_to_the_reference():
    """oracle.calculate_uncertainty / compressed_vocab_size against outputs of the UNMODIFIED reference
    (tests/golden/cuhlm_v1.npz, written by oracle/gen_golden_cuhlm.py)."""
    import os

    import numpy as np
    import torch

    import oracle

    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "cuhlm_v1.npz"))
    for ci in range(3):
        pre = f"unc/{ci}/"
        M, theta_max, draft = z[pre + "params"]
        temps = torch.clamp(torch.from_numpy(z[pre + "u"]) * float(theta_max), min=1e-6)
        got = oracle.calculate_uncertainty(torch.from_numpy(z[pre + "logits"]), temps, torch.from_numpy(z[pre + "noise"]),
                                           int(draft))
        assert got == float(z[pre + "value"][0]), ci
    for ci in range(5):
        pre = f"kstar/{ci}/"
        unc, theta, draft = z[pre + "params"]
        probs = torch.from_numpy(z[pre + "probs"])
        got = oracle.compressed_vocab_size(float(unc), probs, len(probs), float(theta), None if draft < 0 else int(draft))
        assert got == int(z[pre + "k"][0]), (ci, got)


def test_rank_helpers_oracle_is_pinned_to_the_reference():
    """The oracle's restatement of the rank bookkeeping (src/baselines.py:121-168) against outputs of the unmodified
    reference functions (tests/golden/ranks_v1.npz, written by oracle/gen_golden_ranks.py)."""
    import numpy as np
    import os
    import torch
    import oracle as O

    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "ranks_v1.npz"))
    cases = sorted({k.split("/")[1] for k in z.files})
    assert len(cases) == 6
    for ci in cases:
        pre = f"ranks/{ci}/"
        prefix, accepted, k = (int(v) for v in z[pre + "params"])
        vr, it, tr = [], [], []
        O.record_accepted_token_ranks(stage_probs=torch.from_numpy(z[pre + "probs"]), x=torch.from_numpy(z[pre + "x"]),
                                      prefix_len=prefix, accepted_count=accepted, transfer_top_k=k,
                                      vocab_rank_history=vr, in_transfer_topk_history=it, transfer_topk_rank_history=tr)
        assert vr == z[pre + "vocab_rank"].tolist() and it == z[pre + "in_topk"].tolist(), ci
        assert tr == z[pre + "topk_rank"].tolist(), ci


def test_bf16_reference_goldens_vs_fp32_oracle():
    """The canonicalisation "bf16 rows are evaluated in fp32" pinned on the CPU: the fp32 oracle on the bf16 logits of
    tests/golden/bf16_v1.npz agrees with the reference's own bf16 arithmetic to bf16 precision, and in every decision
    outside bf16 rounding of p/q."""
    import torch
    import oracle as O
    from tests.bf16_check import check_against_bf16_goldens

    def run(dl, tl, toks, r, T, top_k, top_p):
        return O.fused_verify(dl.float(), tl.float(), toks, r, torch.ones(toks.shape[0], tl.shape[-1]), T, top_k, top_p)

    check_against_bf16_goldens(run)
