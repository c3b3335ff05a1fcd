"""Torch-CPU restatement of the reference verification hot path (TEST ORACLE).

Every function states the reference lines (under /root/reference) it follows.
The arithmetic is issued as the same sequence of ATen CPU operators the
reference issues, so on tie-free inputs the outputs are bit-identical to the
reference (pinned by tests/test_oracle_golden.py against fixtures produced by
oracle/gen_golden.py from the unmodified reference).

Two things are *defined here* because the reference leaves them to the library
(SURVEY.md facts 4/5):

* ties: equal logits / probabilities are ordered lowest-index-first
  (``torch.sort(stable=True)``); the reference's ``torch.sort``/``torch.topk``
  give an arbitrary order there.
* sampling noise: ``torch.multinomial(p, 1)`` is ``argmax(p / Exp(1))``
  (ATen ``multinomial_out`` fast path); the oracle takes the Exp(1) draws as an
  argument so the CUDA path and the oracle can be fed identical noise.

B > 1: the reference reads row 0 only (src/decoding_ops.py:146); batch parity
is "reference applied to every row on its own" (``fused_verify``).

This module is test infrastructure: product code never imports it.
"""

from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Optional, Tuple

import torch
import torch.nn.functional as F

__all__ = [
    "token_vocab_rank",
    "transfer_topk_rank",
    "record_accepted_token_ranks",
    "filter_logits",
    "norm_logits",
    "sample",
    "race_argmax",
    "max_fn",
    "topk_canonical",
    "rebuild_topk_probs",
    "rebuild_topk_uniform_probs",
    "collect_verification_payload",
    "accept_rows",
    "compute_acceptance",
    "RollbackPlanO",
    "build_rollback_plan",
    "residual_distribution",
    "sample_reject_token",
    "sample_accept_token",
    "apply_top_k_compression",
    "rebuild_full_probs",
    "compress_rebuild_probs",
    "transfer_bytes",
    "fused_verify",
    "finalize_positions",
    "top_p_margin",
    "accept_margin",
    "race_gap",
    "calculate_uncertainty_tokens",
    "calculate_uncertainty",
    "softplus",
    "compressed_vocab_size",
]


# --------------------------------------------------------------------------
# logits -> probabilities
# --------------------------------------------------------------------------
def filter_logits(logits: torch.Tensor, top_k: int = 0, top_p: float = 0.0) -> torch.Tensor:
    """Top-k then top-p (nucleus) masking to -inf.  src/utils.py:650-683.

    top-k keeps every entry >= the k-th largest value (ties survive, :665-666).
    top-p sorts descending, takes softmax+cumsum over the sorted row and drops
    sorted position j when cum[j-1] > top_p (mask shifted right by one, first
    position always kept, :669-681).  The CPU cumsum accumulates in float64 and
    rounds each prefix to the row dtype; the comparison is against top_p rounded
    to the row dtype.  Ties are ordered lowest index first (stable sort).
    """
    if top_k > 0:
        k = min(int(top_k), logits.size(-1))
        kth = torch.topk(logits, k, dim=-1)[0][..., -1, None]
        logits = logits.masked_fill(logits < kth, float("-inf"))
    if top_p > 0.0:
        srt, order = torch.sort(logits, descending=True, dim=-1, stable=True)
        cum = torch.cumsum(F.softmax(srt, dim=-1), dim=-1)
        drop = cum > top_p
        drop[..., 1:] = drop[..., :-1].clone()
        drop[..., 0] = 0
        drop_at_index = drop.scatter(-1, order, drop)
        logits = logits.masked_fill(drop_at_index, float("-inf"))
    return logits


def norm_logits(logits: torch.Tensor, temperature: float, top_k: float, top_p: float) -> torch.Tensor:
    """Temperature / filter / softmax.  src/utils.py:686-709.

    temperature == 0: fp32 one-hot of the (first) argmax (:700-704); otherwise
    ``softmax(filter(logits / T))`` in the input dtype (:706-709).
    """
    if temperature == 0:
        hot = logits.argmax(dim=-1, keepdim=True)
        out = torch.zeros_like(logits)
        out.scatter_(-1, hot, 1)
        return out.float()
    scaled = logits / temperature
    scaled = filter_logits(scaled, top_k=int(top_k), top_p=top_p)
    return F.softmax(scaled, dim=-1)


# --------------------------------------------------------------------------
# sampling
# --------------------------------------------------------------------------
def race_argmax(probs: torch.Tensor, noise: torch.Tensor) -> torch.Tensor:
    """``torch.multinomial(probs, 1)`` with its Exp(1) draws supplied.

    ATen ``multinomial_out`` (n_sample == 1): ``q = exponential_(1);
    argmax(probs / q)``; argmax returns the first maximal index.
    """
    return torch.argmax(probs / noise, dim=-1, keepdim=True)


def sample(probs: torch.Tensor, noise: Optional[torch.Tensor] = None, num_samples: int = 1) -> torch.Tensor:
    """src/utils.py:727-748: fp32, nan/inf -> 0, clamp >= 0, zero-mass rows fall
    back to a one-hot at argmax, renormalise, multinomial."""
    p = probs.float()
    p = torch.nan_to_num(p, nan=0.0, posinf=0.0, neginf=0.0)
    p = p.clamp_min(0.0)
    mass = p.sum(dim=-1, keepdim=True)
    dead = mass.squeeze(-1) <= 0
    if dead.any():
        hot = torch.zeros_like(p[dead])
        hot.scatter_(-1, p[dead].argmax(dim=-1, keepdim=True), 1.0)
        p = p.clone()
        p[dead] = hot
        mass = p.sum(dim=-1, keepdim=True)
    p = p / mass
    if noise is None:
        return torch.multinomial(p, num_samples=num_samples)
    return race_argmax(p, noise.float())


def max_fn(x: torch.Tensor) -> torch.Tensor:
    """norm(max(x, 0)) over dim 1; zero-mass rows become a one-hot at argmax(x)
    (first index on ties).  src/utils.py:787-810."""
    x = torch.nan_to_num(x.float(), nan=0.0, posinf=0.0, neginf=0.0)
    pos = torch.where(x > 0, x, torch.zeros_like(x))
    mass = torch.sum(pos, dim=1, keepdim=True)
    live = mass.squeeze(-1) > 0
    out = torch.zeros_like(pos)
    if live.any():
        out[live] = pos[live] / mass[live]
    if (~live).any():
        hot = torch.zeros_like(pos[~live])
        hot.scatter_(-1, x[~live].argmax(dim=-1, keepdim=True), 1.0)
        out[~live] = hot
    return out


# --------------------------------------------------------------------------
# transfer_top_k compression
# --------------------------------------------------------------------------
def topk_canonical(values: torch.Tensor, k: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """Sorted top-k along the last dim with lowest-index-first ties (the
    reference's ``torch.topk(..., sorted=True)`` leaves tie order open)."""
    srt, order = torch.sort(values, descending=True, dim=-1, stable=True)
    return srt[..., :k], order[..., :k]


def rebuild_topk_probs(probs: torch.Tensor, top_k: Optional[int], strategy: str = "uniform") -> torch.Tensor:
    """Keep the k largest, spread clamp(1-sum,0) over exact zeros, renormalise.
    src/utils.py:751-777.  Identity when k is None / <= 0 / >= V / empty."""
    if strategy != "uniform":
        raise ValueError(f"Unsupported top-k rebuild strategy: {strategy}")
    if top_k is None or top_k <= 0 or probs.numel() == 0 or top_k >= probs.shape[-1]:
        return probs
    vals, idx = topk_canonical(probs, int(top_k))
    kept = torch.zeros_like(probs)
    kept.scatter_(-1, idx, vals)
    kept_sum = kept.sum(dim=-1, keepdim=True)
    tail_mass = (1.0 - kept_sum).clamp_min(0.0)
    holes = kept == 0
    n_holes = holes.sum(dim=-1, keepdim=True)
    fill = torch.where(n_holes > 0, tail_mass / n_holes, torch.zeros_like(tail_mass))
    rebuilt = torch.where(holes, fill, kept)
    total = rebuilt.sum(dim=-1, keepdim=True).clamp_min(1e-12)
    return rebuilt / total


def rebuild_topk_uniform_probs(probs: torch.Tensor, top_k: Optional[int]) -> torch.Tensor:
    """src/utils.py:780-784."""
    return rebuild_topk_probs(probs, top_k, strategy="uniform")


def apply_top_k_compression(probs: Optional[torch.Tensor], k: int) -> torch.Tensor:
    """1-D top-k scatter into zeros, no renormalisation.
    src/communication.py:288-308 (duplicate at :655-675)."""
    if probs is None or probs.numel() == 0:
        return torch.empty(0)
    if k >= len(probs):
        return probs
    vals, idx = topk_canonical(probs, k)
    out = torch.zeros_like(probs)
    out[idx] = vals
    return out


def rebuild_full_probs(compressed: torch.Tensor) -> torch.Tensor:
    """Uniform refill of exact-zero slots with 1 - sum (no clamp, no renorm).
    src/communication.py:310-346."""
    if compressed is None or compressed.numel() == 0:
        return torch.empty(0)
    kept_sum = compressed.sum(dim=-1, keepdim=True)
    tail_mass = 1.0 - kept_sum
    holes = compressed == 0
    n_holes = holes.sum(dim=-1, keepdim=True)
    fill = torch.where(n_holes > 0, tail_mass / n_holes, torch.zeros_like(tail_mass))
    return torch.where(holes, fill, compressed.clone())


def compress_rebuild_probs(probs: torch.Tensor, k: int) -> torch.Tensor:
    """[B,S,V] top-k + uniform refill in one go.  src/communication.py:348-395."""
    if probs is None or probs.numel() == 0:
        return torch.empty(0)
    if probs.dim() != 3:
        raise ValueError(f"probs must be 3-D, got {probs.dim()}-D")
    if k >= probs.shape[-1]:
        return probs
    b, s, v = probs.shape
    flat = probs.reshape(-1, v)
    vals, idx = topk_canonical(flat, k)
    kept = torch.zeros_like(flat)
    kept.scatter_(-1, idx, vals)
    kept_sum = kept.sum(dim=-1, keepdim=True)
    tail_mass = 1.0 - kept_sum
    holes = kept == 0
    n_holes = holes.sum(dim=-1, keepdim=True)
    fill = torch.where(n_holes > 0, tail_mass / n_holes, torch.zeros_like(tail_mass))
    return torch.where(holes, fill, kept).view(b, s, v)


def transfer_bytes(
    tokens: Optional[torch.Tensor],
    prob: Optional[torch.Tensor],
    protocol_overhead_bytes: int = 0,
    is_compressed: bool = False,
    compressed_k: Optional[int] = 300,
) -> int:
    """Byte-accounting rule of ``CommunicationSimulator.transfer``.
    src/communication.py:397-431: compressed payload counts k values per
    sequence position, indices are not counted."""
    token_bytes = 0
    prob_bytes = 0
    if tokens is not None and tokens.numel() > 0:
        token_bytes = tokens.element_size() * tokens.numel()
    if prob is not None and prob.numel() > 0:
        prob_bytes = prob.element_size() * prob.numel()
    total = token_bytes + prob_bytes + protocol_overhead_bytes
    if is_compressed and prob is not None and prob.numel() > 0 and compressed_k is not None:
        seq = prob.shape[1] if prob.dim() == 3 else 1
        total = token_bytes + compressed_k * prob.element_size() * seq + protocol_overhead_bytes
    return int(total)


# --------------------------------------------------------------------------
# verification
# --------------------------------------------------------------------------
def collect_verification_payload(
    prob_history: torch.Tensor, x: torch.Tensor, prefix_len: int, gamma: int
) -> Tuple[torch.Tensor, torch.Tensor]:
    """Draft tokens and q_j(x_j), truncated to the rows/tokens that exist.
    src/decoding_ops.py:21-47."""
    if gamma <= 0:
        return x[:, 0:0], prob_history[:, 0:0, 0]
    rows_left = max(prob_history.shape[1] - (prefix_len - 1), 0)
    toks_left = max(x.shape[1] - prefix_len, 0)
    g = min(gamma, rows_left, toks_left)
    toks = x[:, prefix_len : prefix_len + g]
    if g <= 0:
        return toks, prob_history[:, 0:0, 0]
    rows = prob_history[:, prefix_len - 1 : prefix_len + g - 1, :]
    return toks, torch.gather(rows, 2, toks.unsqueeze(-1)).squeeze(-1)


def accept_rows(
    draft_probs: torch.Tensor, target_probs: torch.Tensor, tokens: torch.Tensor, r: torch.Tensor
) -> Dict[str, torch.Tensor]:
    """Gather, ``r <= p/q`` (raw IEEE division: q=0,p>0 accepts; 0/0 rejects),
    leading-run length per row.  src/decoding_ops.py:127-145 (the reference then
    reads row 0 only, :146)."""
    g = tokens.shape[1]
    idx = tokens.unsqueeze(-1)
    sel_q = torch.gather(draft_probs[:, :g], 2, idx).squeeze(-1)
    sel_p = torch.gather(target_probs[:, :g], 2, idx).squeeze(-1)
    mask = r <= (sel_p / sel_q)
    run, _ = mask.to(torch.int8).cummin(dim=1)
    return {
        "selected_draft_p": sel_q,
        "selected_target_p": sel_p,
        "accept_mask": mask,
        "accepted_len": run.sum(dim=1).to(torch.int64),
    }


def compute_acceptance(
    draft_probs: torch.Tensor,
    target_probs: torch.Tensor,
    tokens: torch.Tensor,
    prefix_len: int,
    r: torch.Tensor,
) -> Dict[str, object]:
    """``compute_acceptance_result`` incl. the row-0 read-out and n.
    src/decoding_ops.py:109-158."""
    g = tokens.shape[1]
    if g <= 0:
        return {
            "accepted_count": 0,
            "n": prefix_len - 1,
            "selected_draft_p": draft_probs[:, 0:0, 0],
            "selected_target_p": target_probs[:, 0:0, 0],
            "accept_mask": torch.zeros((draft_probs.shape[0], 0), dtype=torch.bool),
        }
    out = accept_rows(draft_probs, target_probs, tokens, r)
    count = int(out["accepted_len"][0].item())
    out["accepted_count"] = count
    out["n"] = prefix_len + count - 1
    return out


@dataclass
class RollbackPlanO:
    draft_end_pos: int
    target_end_pos_reject: int
    target_end_pos_accept: int
    all_accepted: bool


def build_rollback_plan(prefix_len: int, gamma: int, n: int) -> RollbackPlanO:
    """src/decoding_ops.py:168-175."""
    return RollbackPlanO(n + 1, n + 1, n + 2, n >= prefix_len + gamma - 1)


def finalize_positions(prefix_len: int, gamma: int, n: int, cache_len_draft: int, cache_len_target: int):
    """Cache lengths after ``apply_rollback`` (src/decoding_ops.py:178-187) with
    the clamp of ``KVCacheModel.rollback`` (src/model_gpu.py:383)."""
    plan = build_rollback_plan(prefix_len, gamma, n)
    d = min(plan.draft_end_pos, cache_len_draft)
    t_end = plan.target_end_pos_accept if plan.all_accepted else plan.target_end_pos_reject
    return d, min(t_end, cache_len_target), plan.all_accepted


def residual_distribution(target_probs: torch.Tensor, draft_probs: torch.Tensor) -> torch.Tensor:
    """src/decoding_ops.py:161-165."""
    return max_fn(target_probs - draft_probs)


def sample_reject_token(target_probs, draft_probs, noise=None) -> torch.Tensor:
    """src/decoding_ops.py:190-206."""
    return sample(residual_distribution(target_probs, draft_probs), noise)


def sample_accept_token(target_next_probs, noise=None) -> torch.Tensor:
    """src/decoding_ops.py:209-220."""
    return sample(target_next_probs, noise)


def fused_verify(
    draft_logits: torch.Tensor,  # [B, >=gamma, Vd]
    target_logits: torch.Tensor,  # [B, gamma+1, Vt]
    tokens: torch.Tensor,  # [B, gamma] int64
    r: torch.Tensor,  # [B, gamma]
    noise: torch.Tensor,  # [B, Vt] Exp(1)
    temperature: float = 1.0,
    top_k: int = 0,
    top_p: float = 0.0,
    draft_temperature: Optional[float] = None,
    draft_top_k: Optional[int] = None,
    draft_top_p: Optional[float] = None,
    inputs_are_probs: bool = False,
) -> Dict[str, torch.Tensor]:
    """One verification step, reference applied to every row on its own:
    norm_logits on both models (src/model_gpu.py:253), accept test + leading
    run (src/decoding_ops.py:127-147), then on the first rejected position the
    residual resample over min(Vd, Vt) entries (:383-390, :190-206) or, when all
    gamma drafts survive, the bonus sample from the last target row (:377-381).
    """
    B, g = tokens.shape
    dT = temperature if draft_temperature is None else draft_temperature
    dk = top_k if draft_top_k is None else draft_top_k
    dp = top_p if draft_top_p is None else draft_top_p
    vlim = min(draft_logits.shape[-1], target_logits.shape[-1])
    Vt = target_logits.shape[-1]
    acc = torch.zeros(B, dtype=torch.int64)
    tok = torch.zeros(B, dtype=torch.int64)
    mask = torch.zeros(B, g, dtype=torch.bool)
    sel_p = torch.zeros(B, g, dtype=torch.float32)
    sel_q = torch.zeros(B, g, dtype=torch.float32)
    # one norm_logits call per model over [B, gamma(+1), V], as the reference issues it (src/model_gpu.py:253)
    if inputs_are_probs:
        q_all, p_all = draft_logits[:, :g], target_logits[:, : g + 1]
    else:
        q_all = norm_logits(draft_logits[:, :g], dT, dk, dp)
        p_all = norm_logits(target_logits[:, : g + 1], temperature, top_k, top_p)
    for b in range(B):
        q, p = q_all[b : b + 1], p_all[b : b + 1]
        if q.shape[-1] != p.shape[-1]:
            # the accept gather happens on draft-vocab indices in both tensors
            out = accept_rows(q[..., :vlim], p[..., :vlim], tokens[b : b + 1], r[b : b + 1])
        else:
            out = accept_rows(q, p, tokens[b : b + 1], r[b : b + 1])
        a = int(out["accepted_len"][0])
        acc[b] = a
        mask[b] = out["accept_mask"][0]
        sel_p[b] = out["selected_target_p"][0].float()
        sel_q[b] = out["selected_draft_p"][0].float()
        if a < g:
            t = sample_reject_token(p[:, a, :vlim], q[:, a, :vlim], noise[b : b + 1, :vlim])
        else:
            t = sample_accept_token(p[:, g, :Vt], noise[b : b + 1, :Vt])
        tok[b] = int(t)
    return {
        "accepted_len": acc,
        "token": tok,
        "accept_mask": mask,
        "selected_target_p": sel_p,
        "selected_draft_p": sel_q,
        "all_accepted": acc == g,
    }


# --------------------------------------------------------------------------
# near-tie diagnostics (float64): how far each discrete decision is from flipping
# --------------------------------------------------------------------------
def top_p_margin(logits: torch.Tensor, temperature: float, top_k: int, top_p: float) -> torch.Tensor:
    """Per row: distance between top_p and the nearest prefix sum of the sorted
    softmax (float64).  A CUDA/CPU difference in exp/sum rounding of ~1e-7 can
    move the nucleus boundary only on rows whose margin is of that order."""
    x = (logits / temperature).double()
    if top_k > 0:
        k = min(int(top_k), x.size(-1))
        kth = torch.topk(x, k, dim=-1)[0][..., -1, None]
        x = x.masked_fill(x < kth, float("-inf"))
    srt, _ = torch.sort(x, descending=True, dim=-1, stable=True)
    cum = torch.cumsum(F.softmax(srt, dim=-1), dim=-1)
    return (cum - float(torch.tensor(top_p, dtype=torch.float32))).abs().min(dim=-1)[0]


def accept_margin(sel_p: torch.Tensor, sel_q: torch.Tensor, r: torch.Tensor) -> torch.Tensor:
    """|p/q - r| / max(r, tiny) per draft position (float64)."""
    ratio = sel_p.double() / sel_q.double()
    return (ratio - r.double()).abs() / r.double().clamp_min(1e-30)


def race_gap(probs: torch.Tensor, noise: torch.Tensor) -> torch.Tensor:
    """Relative gap between the two largest race keys p_i / e_i per row."""
    keys = probs.double() / noise.double()
    top2 = torch.topk(keys, 2, dim=-1)[0]
    return (top2[..., 0] - top2[..., 1]) / top2[..., 0].clamp_min(1e-300)


# --------------------------------------------------------------------------
# CUHLM helpers (CommunicationSimulator)
# --------------------------------------------------------------------------
def calculate_uncertainty_tokens(logits: torch.Tensor, temperatures: torch.Tensor, noise: torch.Tensor) -> torch.Tensor:
    """The M tempered draws of ``calculate_uncertainty`` (src/communication.py:556-570) with the
    temperatures and the multinomial's Exp(1) noise injected: tokens [M]."""
    if logits.dim() > 1:
        logits = logits[0]
    pert = logits.float().unsqueeze(0) / temperatures.float().unsqueeze(1)
    pert = pert - pert.max(dim=1, keepdim=True)[0]
    return race_argmax(torch.softmax(pert, dim=-1), noise)


def calculate_uncertainty(logits, temperatures, noise, draft_token: int) -> float:
    """src/communication.py:541-575: share of the M draws that differ from the draft token."""
    toks = calculate_uncertainty_tokens(logits, temperatures, noise)
    return int((toks != int(draft_token)).sum().item()) / int(toks.numel())


def softplus(z: float, eta: float = 1.0) -> float:
    """src/communication.py:678-679."""
    return float(torch.log(1 + torch.exp(eta * torch.tensor(float(z)))) / eta)


def compressed_vocab_size(uncertainty: float, current_probs: torch.Tensor, vocab_size: int, theta: float = 0.1,
                          draft_token: Optional[int] = None, default_size: int = 300) -> int:
    """``_calculate_compressed_vocab_size`` (src/communication.py:681-755): smallest k whose uniform-refill
    error bound is <= theta.  Same arithmetic as the reference's loops (fp32 ``torch.sum`` of the head,
    float64 accumulation of the tail), the inner Python loop written as one vector expression per k."""
    if current_probs is None or current_probs.numel() == 0:
        return 0
    if len(current_probs) != vocab_size:
        return max(1, min(300, vocab_size // 100))
    a, b = 0.815, -0.066
    beta_d = max(0, min(1, a * uncertainty + b))
    sorted_probs, _ = torch.sort(current_probs, descending=True)
    if draft_token is None or not (0 <= draft_token < len(current_probs)):
        x_d = sorted_probs[0].item()
    else:
        x_d = current_probs[draft_token].item()
    denominator = (1 - x_d) * softplus(-1.0) + x_d * softplus(-beta_d)
    if denominator <= 0:
        return 30
    s64 = sorted_probs.double()
    for k in range(1, vocab_size):
        top_k_sum = torch.sum(sorted_probs[:k]).item()
        residual = 1.0 - top_k_sum
        uniform = 0 if residual <= 0 else residual / (vocab_size - k)
        numerator = float((s64[k:] - uniform).abs().sum())
        if numerator / denominator <= theta:
            return k
    return min(default_size, vocab_size // 100)


# ---------------------------------------------------------------------------------------------------------
# Rank bookkeeping of accepted tokens
def token_vocab_rank(probs: torch.Tensor, token_id: int) -> int:
    """src/baselines.py:121-123: 1 + number of entries greater than the token's probability."""
    return int((probs > probs[..., token_id]).sum().item()) + 1


def transfer_topk_rank(probs: torch.Tensor, token_id: int, transfer_top_k: Optional[int], vocab_rank: int):
    """src/baselines.py:126-141: (in top-k, 1-based position in the sorted top-k); equal values in index order
    (topk_canonical - torch.topk leaves that order unspecified)."""
    V = probs.shape[-1]
    if transfer_top_k is None or transfer_top_k <= 0 or transfer_top_k >= V:
        return True, vocab_rank
    _, idx = topk_canonical(probs.reshape(1, V), min(int(transfer_top_k), V))
    hit = (idx[0] == token_id).nonzero(as_tuple=False)
    if hit.numel() == 0:
        return False, 0
    return True, int(hit[0].item()) + 1


def record_accepted_token_ranks(*, stage_probs, x, prefix_len, accepted_count, transfer_top_k, vocab_rank_history,
                                in_transfer_topk_history, transfer_topk_rank_history) -> None:
    """src/baselines.py:144-168."""
    if stage_probs is None or accepted_count <= 0:
        return
    for i in range(accepted_count):
        token_id = int(x[:, prefix_len + i].item())
        probs = stage_probs[0, prefix_len + i - 1, :]
        vr = token_vocab_rank(probs, token_id)
        ok, tr = transfer_topk_rank(probs, token_id, transfer_top_k, vr)
        vocab_rank_history.append(vr)
        in_transfer_topk_history.append(ok)
        transfer_topk_rank_history.append(tr)
