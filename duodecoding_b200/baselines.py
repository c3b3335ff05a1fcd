"""Mirrors of the rank bookkeeping helpers of /root/reference/src/baselines.py:121-168.

The reference walks the accepted tokens one by one: ``(probs > p).sum().item()`` for the vocabulary rank, a sorted
``torch.topk(probs, transfer_top_k)`` plus ``.nonzero()`` / ``.item()`` for the rank inside the transmitted top-k -
a sort and three host synchronisations per accepted token.  Here one kernel (``duodec_token_ranks``) counts, for
every accepted position at once, the entries above the token's probability and the equal entries at a lower index;
one device-to-host copy brings all counts back.  Equal values rank in index order (torch.topk leaves the order of
equal values unspecified; the oracle pins the same canonical order).
"""

from __future__ import annotations

from typing import List, Optional, Tuple

import torch

from . import _lib, ops


def _rank_counts(rows: torch.Tensor, tokens: torch.Tensor) -> List[Tuple[int, int]]:
    """rows [n, V] (any row stride), tokens [n] int64 on the same device -> [(greater, equal_before)] per row."""
    ops._need_cuda(rows, tokens)
    n, V = rows.shape
    if n == 0:
        return []
    if rows.dtype not in (torch.float32, torch.bfloat16):
        rows = rows.float()
    if rows.stride(-1) != 1:
        rows = rows.contiguous()
    tokens = tokens.to(torch.int64)
    out = torch.empty((n, 2), dtype=torch.int32, device=rows.device)
    with torch.cuda.device(rows.device):
        _lib.check(_lib.lib.duodec_token_ranks(rows.data_ptr(), ops._code(rows), rows.stride(0), n, V, tokens.data_ptr(),
                                               tokens.stride(0), out.data_ptr(), ops._stream()), "duodec_token_ranks")
    return [(int(a), int(b)) for a, b in out.cpu().tolist()]  # the one synchronisation


def _topk_rank(greater: int, before: int, V: int, transfer_top_k: Optional[int]) -> Tuple[bool, int]:
    vocab_rank = greater + 1
    if transfer_top_k is None or transfer_top_k <= 0 or transfer_top_k >= V:
        return True, vocab_rank
    pos = greater + before + 1  # position in the sorted top-k, equal values in index order
    return (True, pos) if pos <= min(transfer_top_k, V) else (False, 0)


def _compute_token_vocab_rank(probs: torch.Tensor, token_id: int) -> int:
    """src/baselines.py:121-123."""
    row = probs.reshape(-1, probs.shape[-1])[:1]
    (greater, _), = _rank_counts(row, torch.tensor([int(token_id)], device=row.device))
    return greater + 1


def _compute_transfer_topk_rank(probs: torch.Tensor, token_id: int, transfer_top_k: Optional[int],
                                vocab_rank: int) -> Tuple[bool, int]:
    """src/baselines.py:126-141."""
    V = probs.shape[-1]
    if transfer_top_k is None or transfer_top_k <= 0 or transfer_top_k >= V:
        return True, vocab_rank
    row = probs.reshape(-1, V)[:1]
    (greater, before), = _rank_counts(row, torch.tensor([int(token_id)], device=row.device))
    return _topk_rank(greater, before, V, transfer_top_k)


def _record_accepted_token_ranks(*, stage_probs: Optional[torch.Tensor], x: torch.Tensor, prefix_len: int,
                                 accepted_count: int, transfer_top_k: Optional[int], vocab_rank_history: List[int],
                                 in_transfer_topk_history: List[bool], transfer_topk_rank_history: List[int]) -> None:
    """src/baselines.py:144-168: ranks of the ``accepted_count`` tokens ``x[:, prefix_len + i]`` in the rows
    ``stage_probs[0, prefix_len + i - 1]`` appended to the three histories - one launch, one read-back."""
    if stage_probs is None or accepted_count <= 0:
        return
    n = int(accepted_count)
    rows = stage_probs[0, prefix_len - 1:prefix_len - 1 + n, :]
    tokens = x[0, prefix_len:prefix_len + n].to(rows.device)
    if rows.shape[0] != n or tokens.shape[0] != n:
        raise IndexError("accepted positions lie outside the probability history / the token sequence")
    V = rows.shape[-1]
    for greater, before in _rank_counts(rows, tokens):
        if greater < 0:
            raise IndexError("accepted token id outside the vocabulary")
        in_topk, rank = _topk_rank(greater, before, V, transfer_top_k)
        vocab_rank_history.append(greater + 1)
        in_transfer_topk_history.append(in_topk)
        transfer_topk_rank_history.append(rank)
