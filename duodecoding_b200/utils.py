"""Logit warp, sampling and top-k rebuild with the reference's names and signatures.

Mirror of the hot-path functions of /root/reference/src/utils.py (file:line in each docstring);
all arithmetic runs in the sm_100a kernels behind :mod:`duodecoding_b200.ops`.  CUDA tensors only.

Randomness: ``sample`` draws its Exp(1) race noise with torch's generator on the device (exactly
the draw ``torch.multinomial`` makes internally); pass ``noise=`` to inject it.
"""

from __future__ import annotations

from typing import Optional

import torch

from . import ops


def top_k_top_p_filter(logits: torch.Tensor, top_k: int = 0, top_p: float = 0.0) -> torch.Tensor:
    """src/utils.py:650-683: logits with everything outside the top-k / nucleus set to -inf.

    The kept set is computed by the norm_logits kernel (ties: lowest index first); entries whose
    probability is positive are the kept ones.
    """
    if (top_k is None or top_k <= 0) and (top_p is None or top_p <= 0.0):
        return logits
    probs = ops.norm_logits(logits, 1.0, int(top_k or 0), float(top_p or 0.0), out_dtype=torch.float32)
    # a kept entry more than ~88 nats below the row maximum underflows to 0 in fp32 and is masked too
    return logits.masked_fill(probs <= 0, float("-inf"))


def norm_logits(logits: torch.Tensor, temperature: float, top_k: float, top_p: float) -> torch.Tensor:
    """src/utils.py:686-709: one-hot argmax (fp32) at T == 0, else softmax(filter(logits / T))."""
    return ops.norm_logits(logits, float(temperature), int(top_k or 0), float(top_p or 0.0))


def sample(probs: torch.Tensor, num_samples: int = 1, noise: Optional[torch.Tensor] = None) -> torch.Tensor:
    """src/utils.py:727-748: one categorical draw per row -> [..., 1] int64 (zero rows -> argmax)."""
    if num_samples != 1:
        raise NotImplementedError("the verification path draws one token per row")
    return ops.sample_rows(probs, noise)


def max_fn(x: torch.Tensor) -> torch.Tensor:
    """src/utils.py:787-810: norm(max(x, 0)); zero-mass rows become one-hot at the first argmax."""
    return ops.max_fn(x)


def rebuild_topk_probs(probs: torch.Tensor, top_k: Optional[int], strategy: str = "uniform") -> torch.Tensor:
    """src/utils.py:751-781: keep the top-k values, spread the rest uniformly, renormalise."""
    if strategy != "uniform":
        raise ValueError(f"unknown rebuild strategy: {strategy}")
    if top_k is None or top_k <= 0 or top_k >= probs.shape[-1]:
        return probs
    return ops.topk_rebuild(probs, int(top_k), 0)


def rebuild_topk_uniform_probs(probs: torch.Tensor, top_k: Optional[int]) -> torch.Tensor:
    """src/utils.py:784."""
    return rebuild_topk_probs(probs, top_k, "uniform")
