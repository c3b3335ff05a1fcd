"""Result / plan records of the verification step.

Mirror of /root/reference/src/decoding_types.py:8-34 (same class and field names, so that the
reference's callers and tests read them unchanged).
"""

from __future__ import annotations

from dataclasses import dataclass

import torch


@dataclass
class VerificationInputs:
    """Rows `[prefix_len-1, max_idx)` of both probability histories plus the drafted ids."""

    draft_probs_batch: torch.Tensor      # [B, actual_gamma, Vd]
    target_probs_batch: torch.Tensor     # [B, actual_gamma, Vt] on the draft device
    draft_tokens: torch.Tensor           # [B, actual_gamma] int64
    draft_token_indices: torch.Tensor    # [B, actual_gamma, 1]
    prefix_len: int
    gamma: int
    actual_gamma: int
    max_idx: int


@dataclass
class AcceptanceResult:
    accepted_count: int                  # leading accepted run of sequence 0 (reference semantics)
    n: int                               # absolute index of the last accepted position
    selected_draft_p: torch.Tensor       # [B, actual_gamma]
    selected_target_p: torch.Tensor      # [B, actual_gamma]
    accept_mask: torch.Tensor            # [B, actual_gamma] bool


@dataclass
class RollbackPlan:
    draft_end_pos: int
    target_end_pos_reject: int
    target_end_pos_accept: int
    all_accepted: bool
