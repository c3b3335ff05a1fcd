"""Draft-proposal helpers (mirror of /root/reference/src/proposal_utils.py)."""

from __future__ import annotations

from typing import Optional

import torch


def proposal_top_k(transfer_top_k: Optional[int]) -> Optional[int]:
    """src/proposal_utils.py:8-11: None / <= 0 disables the compression."""
    return transfer_top_k if transfer_top_k is not None and transfer_top_k > 0 else None


class _StageHistory:
    """[B, P-1+gamma, V] view of ``cat(history[:, :P-1], rebuilt)`` that never materialises the cat.

    The reference copies the whole history every step (src/proposal_utils.py:21-27: L=512, V=32000
    bf16 is 32 MB per step) although verification only reads rows >= P-1; this object serves those
    rows from ``rebuilt`` and older ones from the cache's own buffer."""

    def __init__(self, history: torch.Tensor, start: int, rebuilt: torch.Tensor):
        self.history, self.start, self.rebuilt = history, start, rebuilt
        self.shape = torch.Size((history.shape[0], start + rebuilt.shape[1], history.shape[2]))
        self.device, self.dtype = rebuilt.device, rebuilt.dtype

    def dim(self) -> int:
        return 3

    def materialize(self) -> torch.Tensor:
        return torch.cat((self.history[:, :self.start, :], self.rebuilt), dim=1)

    def __getitem__(self, idx):
        if isinstance(idx, tuple) and len(idx) >= 2 and isinstance(idx[1], slice):
            lo, hi, step = idx[1].indices(self.shape[1])
            if step == 1 and lo >= self.start:  # the rows verification reads
                return self.rebuilt[(idx[0], slice(lo - self.start, hi - self.start)) + tuple(idx[2:])]
        if isinstance(idx, tuple) and len(idx) >= 2 and isinstance(idx[1], int):
            row = idx[1] if idx[1] >= 0 else self.shape[1] + idx[1]
            if row >= self.start:
                return self.rebuilt[(idx[0], row - self.start) + tuple(idx[2:])]
        return self.materialize()[idx]


def build_draft_probs_override(cache, stage_start_len: int, rebuilt_draft_probs: Optional[torch.Tensor],
                               lazy: bool = True) -> Optional[torch.Tensor]:
    """src/proposal_utils.py:14-27.  By default the copy-free view above is returned: every access the
    verification ops make (``.shape[1]``, ``[:, a:b, :]`` with a >= P-1, ``[:, n, :V]``) gives the tensor the
    reference's ``cat`` would give, without copying the history; ``lazy=False`` builds the reference's tensor."""
    if rebuilt_draft_probs is None:
        return None
    if lazy:
        return _StageHistory(cache.prob_history, stage_start_len - 1, rebuilt_draft_probs)
    return torch.cat((cache.prob_history[:, :stage_start_len - 1, :], rebuilt_draft_probs), dim=1)


def stage_prob_history(cache, stage_start_len: int, rebuilt_draft_probs: Optional[torch.Tensor]) -> torch.Tensor:
    """src/proposal_utils.py:30-40."""
    override = build_draft_probs_override(cache, stage_start_len, rebuilt_draft_probs)
    return cache.prob_history if override is None else override
