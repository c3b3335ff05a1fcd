"""Known-answer vectors held by the reference's own unit tests, replayed on the oracle.
Sources (under /root/reference/test): test_temperature_sampling.py,
test_engine_finalize_verification.py, test_decoding_ops_stage_verification.py,
test_decoding_ops_payload.py."""
import torch

import oracle as O


def test_temperature_sharpens():  # test_temperature_sampling.py:24-37
    x = torch.tensor([[0.0, 1.0, 2.0]])
    cold, hot = O.norm_logits(x, 0.5, 0, 0.0), O.norm_logits(x, 2.0, 0, 0.0)
    assert torch.allclose(cold.sum(-1), torch.ones(1)) and torch.allclose(hot.sum(-1), torch.ones(1))
    assert cold[0, 2] > hot[0, 2] and cold[0, 0] < hot[0, 0]


def test_top_k_zeroes_tail():  # :131-140
    p = O.norm_logits(torch.tensor([[0.1, 0.2, 3.0, 2.5]]), 1.0, 2, 0.0)
    assert p[0, 0] == 0 and p[0, 1] == 0 and p[0, 2] > 0 and p[0, 3] > 0
    assert torch.allclose(p.sum(-1), torch.ones(1))


def test_top_p_zeroes_tail():  # :142-151
    p = O.norm_logits(torch.tensor([[3.0, 2.0, 0.5, -1.0]]), 1.0, 0, 0.8)
    assert p[0, 0] > 0 and p[0, 1] > 0 and p[0, 2] == 0 and p[0, 3] == 0


def test_accept_rule_known_answer():  # :79-108
    q = torch.tensor([[[0.6, 0.4], [0.3, 0.7]]])
    p = torch.tensor([[[0.3, 0.7], [0.8, 0.2]]])
    out = O.compute_acceptance(q, p, torch.tensor([[1, 1]]), 1, torch.tensor([[0.5, 0.5]]))
    assert out["accepted_count"] == 1 and out["n"] == 1
    assert out["accept_mask"].tolist() == [[True, False]]
    assert torch.allclose(out["selected_draft_p"], torch.tensor([[0.4, 0.7]]))
    assert torch.allclose(out["selected_target_p"], torch.tensor([[0.7, 0.2]]))


def test_zero_draft_probability_accepts():  # :176-197
    out = O.compute_acceptance(
        torch.tensor([[[1.0, 0.0]]]), torch.tensor([[[0.25, 0.75]]]), torch.tensor([[1]]), 1, torch.tensor([[0.5]])
    )
    assert out["accepted_count"] == 1 and out["n"] == 1 and out["accept_mask"].tolist() == [[True]]


def test_residual_distribution_known_answer():  # :110-129
    res = O.residual_distribution(torch.tensor([[0.5, 0.3, 0.2]]), torch.tensor([[0.2, 0.6, 0.2]]))
    assert torch.allclose(res, torch.tensor([[1.0, 0.0, 0.0]]))


def test_zero_residual_mass_falls_back_to_index_zero():  # :199-208
    p = torch.tensor([[0.2, 0.3, 0.5]])
    tok = O.sample_reject_token(p, p.clone(), torch.ones(1, 3))
    assert int(tok) == 0


def test_rebuild_topk_uniform_known_answer():  # :153-160
    out = O.rebuild_topk_uniform_probs(torch.tensor([[0.5, 0.3, 0.15, 0.05]]), 2)
    assert torch.allclose(out[0, :2], torch.tensor([0.5, 0.3])) and torch.allclose(out[0, 2:], torch.tensor([0.1, 0.1]))
    try:
        O.rebuild_topk_probs(torch.tensor([[0.5, 0.5]]), 1, strategy="tail")  # :170-174
        raise AssertionError("expected ValueError")
    except ValueError:
        pass


def test_rollback_positions():  # test_engine_finalize_verification.py:42-74
    plan = O.build_rollback_plan(2, 2, 1)
    assert (plan.draft_end_pos, plan.target_end_pos_reject, plan.all_accepted) == (2, 2, False)
    plan = O.build_rollback_plan(2, 2, 3)
    assert (plan.draft_end_pos, plan.target_end_pos_accept, plan.all_accepted) == (4, 5, True)
    # test_decoding_ops_stage_verification.py:104-110: n = 0 -> both caches to 1
    plan = O.build_rollback_plan(1, 1, 0)
    assert (plan.draft_end_pos, plan.target_end_pos_reject, plan.all_accepted) == (1, 1, False)
    # KVCacheModel.rollback clamps to the real cache length (test_model_gpu_rollback.py:10-37)
    assert O.finalize_positions(6, 2, 6, 6, 6)[:2] == (6, 6)


def test_payload_truncation_known_answers():  # test_decoding_ops_payload.py
    ph = torch.zeros((1, 1, 4))
    ph[0, 0, 2] = 1.0
    tk, pq = O.collect_verification_payload(ph, torch.tensor([[0, 2, 3]]), 1, 2)
    assert tk.tolist() == [[2]] and pq.tolist() == [[1.0]]
    ph = torch.zeros((1, 3, 4))
    ph[0, 0, 1] = ph[0, 1, 2] = ph[0, 2, 3] = 1.0
    tk, pq = O.collect_verification_payload(ph, torch.tensor([[0, 1, 2]]), 1, 5)
    assert tk.tolist() == [[1, 2]] and pq.tolist() == [[1.0, 1.0]]
    tk, pq = O.collect_verification_payload(torch.zeros((1, 1, 4)), torch.tensor([[0]]), 1, 3)
    assert tuple(tk.shape) == (1, 0) and tuple(pq.shape) == (1, 0)


def test_canonical_ties_lowest_index_first():
    # our canonicalisation (the reference leaves tie order to torch.sort/topk)
    vals, idx = O.topk_canonical(torch.tensor([0.2, 0.4, 0.4, 0.1, 0.4]), 2)
    assert idx.tolist() == [1, 2]
    p = O.norm_logits(torch.tensor([[1.0, 1.0, 1.0, 1.0]]), 1.0, 0, 0.5)
    # cum = .25 .5 .75 1: positions with preceding mass <= .5 survive -> first three
    assert (p[0] > 0).tolist() == [True, True, True, False]
    assert int(O.race_argmax(torch.tensor([[0.5, 0.5]]), torch.ones(1, 2))) == 0
