"""Host-side logic of the reference-named mirror modules (no kernels: runs without a GPU).

Modelled on the reference's own tests: test/test_decoding_ops_payload.py,
test/test_engine_finalize_verification.py, test/test_decoding_ops_stage_verification.py (mocks).
"""
from unittest.mock import patch

import pytest
import torch

import oracle
from duodecoding_b200 import communication, decoding_ops, proposal_utils
from duodecoding_b200.decoding_types import AcceptanceResult, RollbackPlan, VerificationInputs


class FakeCache:
    def __init__(self, probs, vocab_size, device="cpu"):
        self.prob_history = probs.clone()
        self.vocab_size = vocab_size
        self.device = torch.device(device)
        self.rollback_calls = []

    def rollback(self, end_pos):
        self.rollback_calls.append(end_pos)


def test_rollback_plan_matches_oracle():
    for P, g, n in [(5, 4, 4), (5, 4, 6), (5, 4, 8), (1, 1, 0), (3, 0, 2)]:
        plan = decoding_ops.build_rollback_plan(P, g, n)
        want = oracle.build_rollback_plan(P, g, n)
        assert (plan.draft_end_pos, plan.target_end_pos_reject, plan.target_end_pos_accept, plan.all_accepted) == (
            want.draft_end_pos, want.target_end_pos_reject, want.target_end_pos_accept, want.all_accepted)


def test_payload_truncation():  # test/test_decoding_ops_payload.py
    g = torch.Generator().manual_seed(0)
    hist = torch.rand(2, 6, 11, generator=g)
    x = torch.randint(0, 11, (2, 8), generator=g)
    for P, gamma in [(3, 3), (5, 4), (6, 2), (7, 3), (3, 0)]:
        toks, q = decoding_ops.collect_verification_payload(hist, x, P, gamma)
        wt, wq = oracle.collect_verification_payload(hist, x, P, gamma)
        assert torch.equal(toks, wt) and torch.equal(q, wq)


def test_prepare_inputs_views_and_missing_history():
    hist_q, hist_p = torch.rand(1, 7, 9), torch.rand(1, 8, 9)
    x = torch.randint(0, 9, (1, 9))
    d, t = FakeCache(hist_q, 9), FakeCache(hist_p, 9)
    vi = decoding_ops.prepare_verification_inputs(d, t, x, prefix_len=4, gamma=4)
    assert vi.actual_gamma == 4 and vi.max_idx == 7
    assert torch.equal(vi.draft_probs_batch, hist_q[:, 3:7]) and torch.equal(vi.target_probs_batch, hist_p[:, 3:7])
    assert torch.equal(vi.draft_tokens, x[:, 4:8]) and vi.draft_token_indices.shape == (1, 4, 1)
    vi = decoding_ops.prepare_verification_inputs(d, t, x, prefix_len=8, gamma=4)  # nothing left to verify
    assert vi.actual_gamma <= 0 or vi.draft_probs_batch.shape[1] == 0
    d.prob_history = None
    with pytest.raises(ValueError):
        decoding_ops.prepare_verification_inputs(d, t, x, prefix_len=4, gamma=4)
    empty = decoding_ops.compute_acceptance_result(VerificationInputs(
        hist_q[:, 0:0], hist_p[:, 0:0], x[:, 0:0], x[:, 0:0].unsqueeze(-1), 4, 4, 0, 3))
    assert (empty.accepted_count, empty.n) == (0, 3) and empty.accept_mask.shape == (1, 0)


def test_resolve_stage_limits_resample_to_common_vocab():  # test/test_decoding_ops_stage_verification.py:21-110
    q = torch.tensor([[[0.6, 0.4, 0.0, 0.0], [0.6, 0.4, 0.0, 0.0]]])
    p = torch.tensor([[[0.1, 0.2, 0.3, 0.4], [0.1, 0.2, 0.3, 0.4]]])
    proposer, verifier = FakeCache(q, 2), FakeCache(p, 4)
    vi = VerificationInputs(q[:, :1], p[:, :1], torch.tensor([[1]]), torch.tensor([[[1]]]), 1, 1, 1, 1)
    res = AcceptanceResult(0, 0, torch.tensor([[0.4]]), torch.tensor([[0.2]]), torch.tensor([[False]]))
    seen = {}

    def fake_reject(target_probs, draft_probs, output_device=None):
        seen["shapes"] = (tuple(target_probs.shape), tuple(draft_probs.shape))
        return torch.tensor([[0]])

    with patch("duodecoding_b200.decoding_ops.verify_draft_sequence_result", return_value=(vi, res)), \
            patch("duodecoding_b200.decoding_ops.sample_reject_token", side_effect=fake_reject):
        acc, n, t, all_acc = decoding_ops.resolve_stage_verification(
            proposer, verifier, torch.tensor([[5, 1]]), 1, 1, output_device=torch.device("cpu"))
    assert (acc, n, all_acc) == (0, 0, False) and t.tolist() == [[0]]
    assert seen["shapes"] == ((1, 2), (1, 2))
    assert proposer.rollback_calls == [1] and verifier.rollback_calls == [1]


def test_finalize_positions():  # test/test_engine_finalize_verification.py:42-74
    q, p = torch.rand(1, 9, 5), torch.rand(1, 9, 5)
    x = torch.arange(9).reshape(1, 9)
    for n, want_calls in [(5, ([6], [6])), (7, ([8], [9]))]:
        d, t = FakeCache(q, 5), FakeCache(p, 5)
        with patch("duodecoding_b200.decoding_ops.sample_reject_token", return_value=torch.tensor([[3]])) as rj, \
                patch("duodecoding_b200.decoding_ops.sample_accept_token", return_value=torch.tensor([[4]])) as ac:
            out = decoding_ops.finalize_verification(d, t, x, prefix_len=4, gamma=4, n=n)
        assert (d.rollback_calls, t.rollback_calls) == want_calls
        assert out.shape == (1, n + 2) and torch.equal(out[:, :n + 1], x[:, :n + 1])
        assert (rj.called, ac.called) == (n < 7, n >= 7)


def test_lazy_stage_history_reads_like_the_cat():
    cache = FakeCache(torch.rand(2, 10, 7), 7)
    rebuilt = torch.rand(2, 3, 7)
    full = proposal_utils.build_draft_probs_override(cache, 6, rebuilt, lazy=False)
    lazy = proposal_utils.build_draft_probs_override(cache, 6, rebuilt, lazy=True)
    assert full.shape == lazy.shape == (2, 8, 7)
    assert torch.equal(lazy[:, 5:8, :], full[:, 5:8, :]) and torch.equal(lazy[:, 6, :4], full[:, 6, :4])
    assert torch.equal(lazy[:, -1, :], full[:, -1, :]) and torch.equal(lazy[:, 2:7, :], full[:, 2:7, :])
    assert proposal_utils.build_draft_probs_override(cache, 6, None) is None
    assert proposal_utils.proposal_top_k(0) is None and proposal_utils.proposal_top_k(300) == 300


def test_transfer_size_rule_and_identity_paths():
    toks, prob = torch.zeros(1, 4, dtype=torch.int64), torch.zeros(1, 4, 100, dtype=torch.bfloat16)
    assert communication.transfer_bytes(toks, prob, 16) == oracle.transfer_bytes(toks, prob, 16)
    assert communication.transfer_bytes(toks, prob, 16, True, 30) == 32 + 30 * 2 * 4 + 16
    assert communication.transfer_bytes(None, None) == 0
    p1 = torch.rand(10)
    assert communication.apply_top_k_compression(p1, 10) is p1 and communication.apply_top_k_compression(None, 3).numel() == 0
    with pytest.raises(ValueError):
        communication.compress_rebuild_probs(torch.rand(4, 10), 3)


def test_install_rebinds_reference_modules_when_present():
    import os
    import sys

    if not os.path.isdir("/root/reference/src"):
        pytest.skip("the reference tree only exists in the build container")
    sys.path.insert(0, "/root/reference")
    try:
        from duodecoding_b200.install import install, uninstall
        done = install("src")
        import src.decoding_ops as ref_ops
        import src.utils as ref_utils
        assert ref_ops.finalize_verification is decoding_ops.finalize_verification
        assert ref_utils.norm_logits.__module__ == "duodecoding_b200.utils" and len(done["decoding_ops"]) >= 12
        with patch("src.decoding_ops.sample", return_value=torch.tensor([[1]])):  # late binding survives
            pass
        uninstall()
        assert ref_ops.finalize_verification is not decoding_ops.finalize_verification
    finally:
        sys.path.remove("/root/reference")
