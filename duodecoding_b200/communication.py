"""transfer_top_k compression helpers of the reference's CommunicationSimulator.

Mirror of the tensor-valued static helpers of /root/reference/src/communication.py and of the
byte-count rule of ``transfer``; the link-time simulation itself (bandwidth tables, queues,
latency draws) is control plane and out of scope.  CUDA tensors only.
"""

from __future__ import annotations

from typing import Optional

import torch

from . import ops


def apply_top_k_compression(probs: Optional[torch.Tensor], k: int) -> torch.Tensor:
    """src/communication.py:288-308 (dup :655-675): the top-k values of a 1-D distribution scattered
    into zeros, no renormalisation; ``k >= len`` returns the input, empty input an empty tensor."""
    if probs is None or probs.numel() == 0:
        return torch.empty(0)
    if k >= len(probs):
        return probs
    return ops.topk_rebuild(probs.reshape(1, -1), int(k), 1).reshape(probs.shape)


def rebuild_full_probs(compressed_probs: torch.Tensor) -> torch.Tensor:
    """src/communication.py:310-346: spread 1 - sum over the exact-zero slots of every row (no clamp,
    no renormalisation)."""
    if compressed_probs is None or compressed_probs.numel() == 0:
        return torch.empty(0)
    return ops.topk_rebuild(compressed_probs, 0, 3)


def compress_rebuild_probs(probs: torch.Tensor, k: int) -> torch.Tensor:
    """src/communication.py:348-395: top-k + uniform refill in one pass over a [B, S, V] history."""
    if probs is None or probs.numel() == 0:
        return torch.empty(0)
    if probs.dim() != 3:
        raise ValueError(f"probs must have 3 dimensions, got {probs.dim()}")
    if k >= probs.shape[-1]:
        return probs
    return ops.topk_rebuild(probs, int(k), 2)


def transfer_bytes(tokens: Optional[torch.Tensor], prob: Optional[torch.Tensor], protocol_overhead_bytes: int = 0,
                   is_compressed: bool = False, compressed_k: Optional[int] = 300) -> int:
    """The size rule of ``CommunicationSimulator.transfer`` (src/communication.py:397-431): tokens +
    probabilities + overhead; a compressed transfer counts k values per position (indices are not
    counted).  Shapes and element sizes only: no device work."""
    token_bytes = tokens.element_size() * tokens.numel() if tokens is not None and tokens.numel() > 0 else 0
    has_prob = prob is not None and prob.numel() > 0
    prob_bytes = prob.element_size() * prob.numel() if has_prob else 0
    if is_compressed and has_prob and compressed_k is not None:
        prob_bytes = compressed_k * prob.element_size() * (prob.shape[1] if prob.dim() == 3 else 1)
    return token_bytes + prob_bytes + protocol_overhead_bytes


class CommunicationSimulator:
    """Only the static tensor helpers of the reference class (same names, so that
    ``CommunicationSimulator._apply_top_k_compression`` etc. resolve); see the module docstring."""

    _apply_top_k_compression = staticmethod(apply_top_k_compression)
    rebuild_full_probs = staticmethod(rebuild_full_probs)
    compress_rebuild_probs = staticmethod(compress_rebuild_probs)
