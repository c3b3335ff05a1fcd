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


# ------------------------------------------------------------------------------------------------ CUHLM
def uncertainty_tokens(logits: torch.Tensor, temperatures: torch.Tensor, noise: torch.Tensor) -> torch.Tensor:
    """One draw from softmax(logits / T_m) per temperature (``argmax p_i / noise[m, i]``): tokens [M] int64."""
    from . import _lib
    ops._need_cuda(logits, temperatures, noise)
    row = logits[0] if logits.dim() > 1 else logits
    row = row.contiguous()
    if row.dtype not in (torch.float32, torch.bfloat16):
        row = row.float()
    temps = temperatures.to(torch.float32).contiguous()
    nz = noise.to(torch.float32)
    if nz.stride(-1) != 1:
        nz = nz.contiguous()
    M, V = temps.numel(), row.numel()
    out = torch.empty((M,), dtype=torch.int64, device=row.device)
    if M:
        with torch.cuda.device(row.device):
            _lib.check(_lib.lib.duodec_uncertainty_tokens(row.data_ptr(), ops._code(row), V, temps.data_ptr(),
                                                          nz.data_ptr(), nz.stride(0), M, out.data_ptr(),
                                                          ops._stream()), "duodec_uncertainty_tokens")
    return out


def calculate_uncertainty(logits: Optional[torch.Tensor], M: int = 20, theta_max: float = 2.0,
                          draft_token: Optional[int] = None) -> float:
    """src/communication.py:541-575: share of M tempered multinomial draws that differ from the draft token.
    One kernel over the row (M blocks) instead of an [M, V] tensor and five passes over it."""
    if logits is None or logits.numel() == 0:
        return 1.0
    row = logits[0] if logits.dim() > 1 else logits
    if draft_token is None:
        draft_token = int(torch.argmax(row).item())
    temps = torch.clamp(torch.rand(M, device=row.device) * theta_max, min=1e-6)
    noise = ops.draw_noise(M, row.numel(), row.device)
    toks = uncertainty_tokens(row, temps, noise)
    return int((toks != int(draft_token)).sum().item()) / M


def _softplus(z: float, eta: float = 1.0) -> float:
    return float(torch.log(1 + torch.exp(eta * torch.tensor(float(z)))) / eta)


def compressed_vocab_size(uncertainty: float, current_probs: Optional[torch.Tensor], vocab_size: int,
                          theta: float = 0.1, draft_token: Optional[int] = None, default_size: int = 300) -> int:
    """``CUHLM._calculate_compressed_vocab_size`` (src/communication.py:681-755): the smallest k whose uniform-refill
    error bound is <= theta.  The O(V^2) Python search becomes one kernel: every k is a thread, the tail sum
    comes from float64 prefix sums and one binary search (sorted descending => a single crossover)."""
    from . import _lib
    if current_probs is None or current_probs.numel() == 0:
        return 0
    if len(current_probs) != vocab_size:
        return max(1, min(300, vocab_size // 100))
    ops._need_cuda(current_probs)
    beta_d = max(0, min(1, 0.815 * uncertainty - 0.066))
    sorted_probs, _ = torch.sort(current_probs.float(), descending=True)
    if draft_token is None or not (0 <= draft_token < len(current_probs)):
        x_d = float(sorted_probs[0].item())
    else:
        x_d = float(current_probs[draft_token].item())
    denominator = (1 - x_d) * _softplus(-1.0) + x_d * _softplus(-beta_d)
    if denominator <= 0:
        return 30
    prefix = torch.cumsum(sorted_probs.double(), 0)
    k_out = torch.empty((1,), dtype=torch.int32, device=current_probs.device)
    with torch.cuda.device(current_probs.device):
        _lib.check(_lib.lib.duodec_compressed_vocab_size(sorted_probs.data_ptr(), prefix.data_ptr(), int(vocab_size),
                                                         float(denominator), float(theta), k_out.data_ptr(),
                                                         ops._stream()), "duodec_compressed_vocab_size")
    k = int(k_out.item())
    return k if k < vocab_size else min(default_size, vocab_size // 100)


class CUHLM(CommunicationSimulator):
    """The tensor helpers of the reference's ``CUHLM`` (src/communication.py:500-791) under the same names."""

    DEFAULT_COMPRESSED_VOCAB_SIZE = 300
    calculate_uncertainty = staticmethod(calculate_uncertainty)
    softplus = staticmethod(_softplus)

    def __init__(self, vocab_size: int = 32000, uncertainty_threshold: float = 0.8):
        self.vocab_size, self.uncertainty_threshold = int(vocab_size), uncertainty_threshold

    def _calculate_compressed_vocab_size(self, uncertainty: float, current_probs: torch.Tensor, theta: float = 0.1,
                                         draft_token: Optional[int] = None) -> int:
        return compressed_vocab_size(uncertainty, current_probs, self.vocab_size, theta, draft_token,
                                     self.DEFAULT_COMPRESSED_VOCAB_SIZE)
