"""Verification operators with the reference's names, signatures and late binding.

Mirror of /root/reference/src/decoding_ops.py (file:line in each docstring).  The tensor work is
done by the sm_100a kernels behind :mod:`duodecoding_b200.ops`; what stays in Python is the
slicing of history views and the Python ints the reference's callers read.  Internal calls go
through this module's globals, so ``unittest.mock.patch("...decoding_ops.sample")`` and friends
keep working like they do on the reference.

``KVCacheModel`` below means anything with the reference's cache surface: ``prob_history``,
``vocab_size``, ``device`` and ``rollback(end_pos)`` (src/model_gpu.py).

Fused route.  When both caches are :class:`duodecoding_b200.model_gpu.KVCacheModel` objects (they
keep LOGIT rows), no ``draft_probs_override`` is given and nobody has patched this module's
``sample`` / ``max_fn`` / ``sample_reject_token`` / ``sample_accept_token``, then
``resolve_stage_verification`` and the ``verify_draft_sequence`` -> ``finalize_verification`` pair
run ONE ``duodec_verify_fused`` call on the logits rows (softmax + filter of the gamma draft rows and
gamma+1 target rows, accept test, leading run, residual / bonus draw) instead of materialising
probability rows and walking them with four operators.  The reference's batch rule is kept: the
accepted count of sequence 0 positions every sequence (``row0_rule``).  Anything else - overrides,
top-k transfer compression, patched samplers, comm accounting that needs the draft rows, plain
tensors as histories - takes the probability-space operators below, exactly as before.
"""

from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import ops
from .decoding_types import AcceptanceResult, RollbackPlan, VerificationInputs
from .utils import max_fn, rebuild_topk_uniform_probs, sample  # noqa: F401  (patch points)


def collect_verification_payload(prob_history: torch.Tensor, x: torch.Tensor, prefix_len: int,
                                 gamma: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """src/decoding_ops.py:21-47: drafted ids x[:, P:P+g'] and q_j(x_j); g' is cut to what exists."""
    steps = 0
    if gamma > 0:
        steps = min(gamma, max(prob_history.shape[1] - (prefix_len - 1), 0), max(x.shape[1] - prefix_len, 0))
    tokens = x[:, prefix_len:prefix_len + steps] if gamma > 0 else x[:, 0:0]
    if steps <= 0:
        return tokens, prob_history[:, 0:0, 0]
    rows = prob_history[:, prefix_len - 1:prefix_len - 1 + steps, :]
    return tokens, torch.gather(rows, 2, tokens.unsqueeze(-1)).squeeze(-1)


def prepare_verification_inputs(draft_model_cache, target_model_cache, x: torch.Tensor, prefix_len: int,
                                gamma: int,
                                draft_probs_override: Optional[torch.Tensor] = None) -> VerificationInputs:
    """src/decoding_ops.py:50-106: views of rows [P-1, max_idx) of both histories (target rows are
    moved to the draft device), ValueError when a history is missing."""
    q_hist = draft_probs_override if draft_probs_override is not None else draft_model_cache.prob_history
    p_hist = target_model_cache.prob_history
    if q_hist is None or p_hist is None:
        raise ValueError("Probability history is not initialized for verification")
    dev = draft_model_cache.device
    lo = prefix_len - 1
    max_idx = min(prefix_len + gamma - 1, q_hist.shape[1], p_hist.shape[1])
    steps = max(max_idx - lo, 0)
    if steps == 0:
        tokens = x[:, 0:0]
        q_rows, p_rows = q_hist[:, 0:0, :], p_hist[:, 0:0, :].to(dev)
    else:
        tokens = x[:, prefix_len:prefix_len + steps]
        q_rows, p_rows = q_hist[:, lo:max_idx, :], p_hist[:, lo:max_idx, :].to(dev)
    return VerificationInputs(draft_probs_batch=q_rows, target_probs_batch=p_rows, draft_tokens=tokens,
                              draft_token_indices=tokens.unsqueeze(-1), prefix_len=prefix_len, gamma=gamma,
                              actual_gamma=steps, max_idx=max_idx)


def compute_acceptance_result(verification_inputs: VerificationInputs, *,
                              r: Optional[torch.Tensor] = None) -> AcceptanceResult:
    """src/decoding_ops.py:109-158: gather p, q at the drafted ids, ``r <= p / q`` (raw IEEE
    division: q = 0 < p accepts, 0/0 rejects), leading run; the count is sequence 0's, as in the
    reference (``continuous_accept[0]``).  One kernel + one device->host read."""
    vi = verification_inputs
    q_rows, p_rows = vi.draft_probs_batch, vi.target_probs_batch
    if vi.actual_gamma <= 0:
        return AcceptanceResult(accepted_count=0, n=vi.prefix_len - 1, selected_draft_p=q_rows[:, 0:0, 0],
                                selected_target_p=p_rows[:, 0:0, 0],
                                accept_mask=torch.zeros((q_rows.shape[0], 0), dtype=torch.bool, device=q_rows.device))
    if r is None:
        r = torch.rand((q_rows.shape[0], vi.actual_gamma), device=q_rows.device)
    res = ops.accept_probs(q_rows, p_rows, vi.draft_tokens, r)
    accepted = int(res["accepted_len"][0].item())
    return AcceptanceResult(accepted_count=accepted, n=int(vi.prefix_len + accepted - 1),
                            selected_draft_p=res["selected_draft_p"], selected_target_p=res["selected_target_p"],
                            accept_mask=res["accept_mask"])


def compute_residual_distribution(target_probs: torch.Tensor, draft_probs: torch.Tensor) -> torch.Tensor:
    """src/decoding_ops.py:161-165: max_fn(p - q) (the subtraction is fused into the kernel)."""
    return max_fn(target_probs.float() - draft_probs.float())


def build_rollback_plan(prefix_len: int, gamma: int, n: int) -> RollbackPlan:
    """src/decoding_ops.py:168-175."""
    return RollbackPlan(draft_end_pos=n + 1, target_end_pos_reject=n + 1, target_end_pos_accept=n + 2,
                        all_accepted=n >= prefix_len + gamma - 1)


def apply_rollback(draft_model_cache, target_model_cache, rollback_plan: RollbackPlan) -> None:
    """src/decoding_ops.py:178-187."""
    draft_model_cache.rollback(rollback_plan.draft_end_pos)
    target_model_cache.rollback(
        rollback_plan.target_end_pos_accept if rollback_plan.all_accepted else rollback_plan.target_end_pos_reject)

_ORIG_SAMPLE, _ORIG_MAX_FN = sample, max_fn


def _unpatched() -> bool:
    """True while nobody has replaced this module's ``sample`` / ``max_fn`` (mock.patch points)."""
    g = globals()
    return g["sample"] is _ORIG_SAMPLE and g["max_fn"] is _ORIG_MAX_FN


def _fused_ready(draft_cache, target_cache, draft_probs_override) -> bool:
    """Both caches keep logit rows the fused kernel can read, and nothing asks for the probability flow."""
    g = globals()
    if draft_probs_override is not None or not _unpatched():
        return False
    if g["sample_reject_token"] is not _ORIG_REJECT or g["sample_accept_token"] is not _ORIG_ACCEPT:
        return False
    if not (hasattr(draft_cache, "logits_rows") and hasattr(target_cache, "logits_rows")):
        return False
    if draft_cache.vocab_size != target_cache.vocab_size:
        return False
    dT, tT = float(draft_cache._temperature), float(target_cache._temperature)
    return (dT == 0.0) == (tT == 0.0)


def _fused_verify(draft_cache, target_cache, x: torch.Tensor, prefix_len: int, gamma: int,
                  r: Optional[torch.Tensor] = None, noise: Optional[torch.Tensor] = None):
    """One duodec_verify_fused call over the logit rows [P-1, P-1+g') of the draft and [P-1, P-1+g'] of the
    target.  Returns (accepted_count, n, token [B,1], g') or None when the rows are not there (then the
    caller takes the probability-space flow)."""
    lo = prefix_len - 1
    q_len, p_len = draft_cache._current_seq_len, target_cache._current_seq_len
    steps = min(prefix_len + gamma - 1, q_len, p_len) - lo
    if gamma <= 0 or steps <= 0 or p_len != lo + steps + 1 or x.shape[1] < prefix_len + steps:
        return None  # the bonus row must be the target's newest row (reference: prob_history[:, -1])
    dl, tl = draft_cache.logits_rows(lo, lo + steps), target_cache.logits_rows(lo, lo + steps + 1)
    if dl is None or tl is None or not dl.is_cuda or dl.device != tl.device or dl.dtype != tl.dtype:
        return None
    if dl.shape[-1] != tl.shape[-1] or dl.shape[0] != tl.shape[0]:
        return None
    B, V, dev = dl.shape[0], dl.shape[-1], dl.device
    tokens = x[:, prefix_len:prefix_len + steps].to(dev)
    if r is None:
        r = torch.rand((B, steps), device=dev)
    # the Exp(1) race noise and the output tensors live with the target cache: no V-sized allocation per step
    buf = getattr(target_cache, "_fused_buf", None)
    if buf is None or buf[0] != (B, steps, V, dev):
        buf = ((B, steps, V, dev), torch.empty((B, V), dtype=torch.float32, device=dev), None)
    if noise is None:
        noise = buf[1].exponential_(1.0)
    out = ops.verify_fused(dl, tl, tokens, r, noise, float(target_cache._temperature),
                           int(target_cache._top_k or 0), float(target_cache._top_p or 0.0),
                           float(draft_cache._temperature), int(draft_cache._top_k or 0),
                           float(draft_cache._top_p or 0.0), out=buf[2], row0_rule=True)
    try:
        target_cache._fused_buf = (buf[0], buf[1], out)
    except AttributeError:
        pass
    accepted = int(out["accepted_len"][0].item())  # the one device->host read of the step (reference: .item())
    return accepted, int(prefix_len + accepted - 1), out["token"].unsqueeze(1).clone(), steps


def sample_reject_token(target_probs: torch.Tensor, draft_probs: torch.Tensor,
                        output_device: Optional[torch.device] = None) -> torch.Tensor:
    """src/decoding_ops.py:190-206: sample(max_fn(p - q)) -> [B, 1] int64.  One fused kernel
    (residual mass, race arg-max, zero-mass fallback) unless ``sample`` / ``max_fn`` are patched."""
    if target_probs.device != draft_probs.device:
        target_probs = target_probs.to(draft_probs.device)
    if _unpatched():
        token = ops.residual_sample(target_probs, draft_probs)
    else:
        token = sample(compute_residual_distribution(target_probs, draft_probs))
    if output_device is not None and token.device != output_device:
        token = token.to(output_device)
    return token


def sample_accept_token(target_next_probs: torch.Tensor,
                        output_device: Optional[torch.device] = None) -> torch.Tensor:
    """src/decoding_ops.py:209-220: the bonus token sample(p_next) -> [B, 1] int64."""
    token = sample(target_next_probs)
    if output_device is not None and token.device != output_device:
        token = token.to(output_device)
    return token


_ORIG_REJECT, _ORIG_ACCEPT = sample_reject_token, sample_accept_token


def verify_draft_sequence_result(draft_model_cache, target_model_cache, x: torch.Tensor, prefix_len: int, gamma: int,
                                 *, draft_probs_override: Optional[torch.Tensor] = None,
                                 r: Optional[torch.Tensor] = None) -> Tuple[VerificationInputs, AcceptanceResult]:
    """src/decoding_ops.py:331-350."""
    vi = prepare_verification_inputs(draft_model_cache=draft_model_cache, target_model_cache=target_model_cache,
                                     x=x, prefix_len=prefix_len, gamma=gamma,
                                     draft_probs_override=draft_probs_override)
    return vi, compute_acceptance_result(vi, r=r)


_ORIG_VDSR = verify_draft_sequence_result


def verify_draft_sequence(draft_model_cache, target_model_cache, x: torch.Tensor, prefix_len: int, gamma: int,
                          comm_simulator=None, comm_link: str = "edge_cloud", transfer_mode: str = "serial",
                          send_reject_message: bool = True, draft_probs_override: Optional[torch.Tensor] = None,
                          decoding_metrics=None) -> Tuple[int, int]:
    """src/decoding_ops.py:223-328: (accepted_count, n).  ``comm_simulator`` (the reference's link-time
    accounting object, control plane and not part of this package) is driven exactly as the reference
    drives it: ``transfer(tokens, q(tokens), link)`` before the test in "batch_before" mode,
    ``send_reject_message(link)`` after a rejection, one ``transfer(token, q_row, link)`` per
    transmitted position in "serial" mode."""
    needs_rows = comm_simulator is not None and transfer_mode in ("batch_before", "serial")
    if not needs_rows and _fused_ready(draft_model_cache, target_model_cache, draft_probs_override):
        fv = _fused_verify(draft_model_cache, target_model_cache, x, prefix_len, gamma)
        if fv is not None:
            accepted, n, token, steps = fv
            if accepted < steps and send_reject_message and comm_simulator:
                comm_simulator.send_reject_message(comm_link)
            if decoding_metrics is not None:
                decoding_metrics["draft_generated_tokens"] += gamma
                decoding_metrics["draft_accepted_tokens"] += int(n - prefix_len + 1)
            if steps == gamma:  # finalize_verification(n) picks the token up instead of drawing again
                target_model_cache._fused_token = (n, prefix_len, gamma, token)
            return accepted, n
    vi = prepare_verification_inputs(draft_model_cache=draft_model_cache, target_model_cache=target_model_cache,
                                     x=x, prefix_len=prefix_len, gamma=gamma,
                                     draft_probs_override=draft_probs_override)
    if vi.actual_gamma <= 0:  # src/decoding_ops.py: nothing to verify, no traffic, no metrics
        return 0, prefix_len - 1
    if transfer_mode == "batch_before" and comm_simulator is not None:
        comm_simulator.transfer(vi.draft_tokens,
                                torch.gather(vi.draft_probs_batch, 2, vi.draft_token_indices).squeeze(-1), comm_link)
    r = torch.rand((x.shape[0], vi.actual_gamma), device=draft_model_cache.device)
    res = compute_acceptance_result(vi, r=r)
    rejected = res.accepted_count < vi.actual_gamma
    if rejected and send_reject_message and comm_simulator:
        comm_simulator.send_reject_message(comm_link)
    if transfer_mode == "serial" and comm_simulator is not None:
        for i in range(res.accepted_count + (1 if rejected else 0)):
            comm_simulator.transfer(vi.draft_token_indices[0, i, 0], vi.draft_probs_batch[:, i, :].squeeze(0),
                                    comm_link)
    if decoding_metrics is not None:
        decoding_metrics["draft_generated_tokens"] += gamma
        decoding_metrics["draft_accepted_tokens"] += int(res.n - prefix_len + 1)
    return res.accepted_count, int(res.n)


def resolve_stage_verification(proposer_cache, verifier_cache, x: torch.Tensor, prefix_len: int, gamma: int, *,
                               output_device: torch.device,
                               draft_probs_override: Optional[torch.Tensor] = None
                               ) -> Tuple[int, int, torch.Tensor, bool]:
    """src/decoding_ops.py:353-397: verify -> plan -> resample (vocab-limited) or bonus -> rollback."""
    if globals()["verify_draft_sequence_result"] is _ORIG_VDSR and _fused_ready(proposer_cache, verifier_cache,
                                                                                draft_probs_override):
        fv = _fused_verify(proposer_cache, verifier_cache, x, prefix_len, gamma)
        if fv is not None:
            accepted, n, token, steps = fv
            plan = build_rollback_plan(prefix_len, steps, n)
            apply_rollback(proposer_cache, verifier_cache, plan)
            if output_device is not None and token.device != output_device:
                token = token.to(output_device)
            return accepted, n, token, plan.all_accepted
    vocab_limit = min(proposer_cache.vocab_size, verifier_cache.vocab_size)
    vi, res = verify_draft_sequence_result(draft_model_cache=proposer_cache, target_model_cache=verifier_cache, x=x,
                                           prefix_len=prefix_len, gamma=gamma,
                                           draft_probs_override=draft_probs_override)
    plan = build_rollback_plan(prefix_len, vi.actual_gamma, res.n)
    if plan.all_accepted:
        t = sample_accept_token(verifier_cache.prob_history[:, -1, :verifier_cache.vocab_size],
                                output_device=output_device)
    else:
        off = res.n - (prefix_len - 1)
        t = sample_reject_token(vi.target_probs_batch[:, off, :vocab_limit], vi.draft_probs_batch[:, off, :vocab_limit],
                                output_device=output_device)
    apply_rollback(proposer_cache, verifier_cache, plan)
    return res.accepted_count, res.n, t, plan.all_accepted


def finalize_verification(approx_model_cache, target_model_cache, x: torch.Tensor, prefix_len: int, gamma: int, n: int,
                          draft_probs_override: Optional[torch.Tensor] = None) -> torch.Tensor:
    """src/decoding_ops.py:400-442: rollback the draft cache, resample at absolute row n or take the
    bonus at row -1, rollback the target cache, return cat(x[:, :n+1], t)."""
    prefix = x[:, :n + 1]
    plan = build_rollback_plan(prefix_len, gamma, n)
    pend = getattr(target_model_cache, "_fused_token", None)
    if pend is not None and draft_probs_override is None and pend[:3] == (n, prefix_len, gamma) and \
            _fused_ready(approx_model_cache, target_model_cache, None):
        target_model_cache._fused_token = None  # the token verify_draft_sequence drew with this very n
        approx_model_cache.rollback(plan.draft_end_pos)
        target_model_cache.rollback(plan.target_end_pos_accept if plan.all_accepted else plan.target_end_pos_reject)
        return torch.cat((prefix, pend[3].to(prefix.device)), dim=1)
    approx_model_cache.rollback(plan.draft_end_pos)
    q_hist = draft_probs_override if draft_probs_override is not None else approx_model_cache.prob_history
    p_hist = target_model_cache.prob_history
    if plan.all_accepted:
        t = sample_accept_token(p_hist[:, -1, :target_model_cache.vocab_size], output_device=prefix.device)
        target_model_cache.rollback(plan.target_end_pos_accept)
    else:
        t = sample_reject_token(p_hist[:, n, :target_model_cache.vocab_size],
                                q_hist[:, n, :approx_model_cache.vocab_size], output_device=prefix.device)
        target_model_cache.rollback(plan.target_end_pos_reject)
    return torch.cat((prefix, t), dim=1)
