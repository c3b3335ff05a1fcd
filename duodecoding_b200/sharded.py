"""Vocabulary-sharded verification step (BASELINE.json config 5: V = 151936 over 1/2/4/8 GPUs).

One process per GPU (``torch.distributed``, NCCL over NVLink/NVSwitch).  Rank ``g`` holds the
contiguous vocabulary slice ``[v_begin, v_end)`` of every draft / target row - the layout a
tensor-parallel LM head leaves behind - and the step needs two small exchanges:

    stats    local (max, sum exp, drafted-token logit) per row            kernel  duodec_shard_stats
    gather   all ranks' stats: 16 B per row and rank                      all_gather_into_tensor
    accept   merge, p(x)/q(x), r <= p/q, leading run, sampling plan       kernel  duodec_shard_accept
    sample   local residual mass + best race key + zero-mass fallback     kernel  duodec_shard_sample
    gather   all ranks' partials: 24 B per sequence and rank              all_gather_into_tensor
    finish   token of every sequence                                      kernel  duodec_shard_finish

Every rank ends with the same accept masks, lengths and tokens.  The kernels are reached through a
``backend`` object; the product backend is the C ABI (CUDA only - there is no CPU path).  Tests
inject a CPU restatement from ``oracle/`` to exercise the partition / exchange / merge logic under
``gloo`` without a GPU.
"""

from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch
import torch.distributed as dist

ALIGN = 8  # slice boundaries are multiples of 8 entries: 16-byte aligned bf16 rows


def shard_bounds(V: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, ALIGN-aligned slices that cover [0, V); the last rank takes the remainder."""
    per = -(-V // world)
    per = -(-per // ALIGN) * ALIGN
    lo = min(rank * per, V)
    return lo, min(lo + per, V)


class _CudaBackend:
    """The C ABI (include/duodec.h, duodec_shard_*)."""

    def __init__(self):
        from . import _lib, ops
        self.lib, self.ops, self.check = _lib.lib, ops, _lib.check
        self.PLAN, self.PARTIAL = 32, 24

    def _rows(self, t):
        return (t.data_ptr(), t.stride(0), t.stride(1)) if t is not None and t.numel() else (None, 0, 0)

    def row_stats(self, draft, target, tokens, v_begin, t_temp, d_temp):
        B, gamma = tokens.shape
        self.ops._need_cuda(draft, target, tokens)
        stats = torch.empty((B * (2 * gamma + 1), 4), dtype=torch.float32, device=target.device)
        dp, dsb, dss = self._rows(draft)
        tp, tsb, tss = self._rows(target)
        with torch.cuda.device(target.device):
            self.check(self.lib.duodec_shard_stats(dp, tp, self.ops._code(target), dsb, dss, tsb, tss, B, gamma,
                                                   target.shape[-1], int(v_begin), tokens.data_ptr(),
                                                   tokens.stride(0) if gamma else 0, float(t_temp), float(d_temp),
                                                   stats.data_ptr(), self.ops._stream()), "duodec_shard_stats")
        return stats

    def accept(self, gathered, B, gamma, r):
        dev = gathered.device
        out = {
            "accept_mask": torch.empty((B, gamma), dtype=torch.bool, device=dev),
            "selected_target_p": torch.empty((B, gamma), dtype=torch.float32, device=dev),
            "selected_draft_p": torch.empty((B, gamma), dtype=torch.float32, device=dev),
            "accepted_len": torch.empty((B,), dtype=torch.int64, device=dev),
            "all_accepted": torch.empty((B,), dtype=torch.bool, device=dev),
        }
        plan = torch.empty((B, self.PLAN), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            self.check(self.lib.duodec_shard_accept(gathered.data_ptr(), gathered.shape[0], B, gamma, r.data_ptr(),
                                                    r.stride(0) if gamma else 0, out["accept_mask"].data_ptr(),
                                                    out["selected_target_p"].data_ptr(),
                                                    out["selected_draft_p"].data_ptr(), out["accepted_len"].data_ptr(),
                                                    out["all_accepted"].data_ptr(), plan.data_ptr(),
                                                    self.ops._stream()), "duodec_shard_accept")
        return out, plan

    def sample_partial(self, draft, target, plan, noise, v_begin, t_temp, d_temp, gamma):
        B = target.shape[0]
        nslice = max(1, min(-(-target.shape[-1] // 4096), -(-592 // max(B, 1))))  # >= 4 blocks per SM in total
        partial = torch.empty((B, nslice, self.PARTIAL), dtype=torch.uint8, device=target.device)
        dp, dsb, dss = self._rows(draft)
        tp, tsb, tss = self._rows(target)
        with torch.cuda.device(target.device):
            self.check(self.lib.duodec_shard_sample(dp, tp, self.ops._code(target), dsb, dss, tsb, tss, B, gamma,
                                                    target.shape[-1], int(v_begin), float(t_temp), float(d_temp),
                                                    plan.data_ptr(), noise.data_ptr(), noise.stride(0), nslice,
                                                    partial.data_ptr(), self.ops._stream()), "duodec_shard_sample")
        return partial

    def finish(self, gathered, B):
        token = torch.empty((B,), dtype=torch.int64, device=gathered.device)
        with torch.cuda.device(gathered.device):
            self.check(self.lib.duodec_shard_finish(gathered.data_ptr(), gathered.shape[0], B, gathered.shape[2],
                                                    token.data_ptr(),
                                                    self.ops._stream()), "duodec_shard_finish")
        return token


class ShardedVerifier:
    """Verification step over vocabulary slices; call :meth:`verify` with THIS rank's slices."""

    def __init__(self, V: int, rank: Optional[int] = None, world: Optional[int] = None, group=None, backend=None):
        self.group = group
        self.world = dist.get_world_size(group) if world is None else int(world)
        self.rank = dist.get_rank(group) if rank is None else int(rank)
        self.V = int(V)
        self.v_begin, self.v_end = shard_bounds(self.V, self.rank, self.world)
        if backend is None:
            if not torch.cuda.is_available():
                raise RuntimeError("ShardedVerifier needs a CUDA device (duodecoding_b200 has no CPU path)")
            backend = _CudaBackend()
        self.backend = backend
        self.collectives = 0

    def slice_of(self, full: torch.Tensor) -> torch.Tensor:
        """This rank's columns of a [..., V] tensor (what a sharded LM head would have produced)."""
        return full[..., self.v_begin:self.v_end]

    def _all_gather(self, local: torch.Tensor) -> torch.Tensor:
        if self.world == 1:
            return local.unsqueeze(0)
        local = local.contiguous()
        out = torch.empty((self.world * local.shape[0],) + tuple(local.shape[1:]), dtype=local.dtype,
                          device=local.device)
        dist.all_gather_into_tensor(out, local, group=self.group)  # rank-major concatenation along dim 0
        self.collectives += 1
        return out.view((self.world,) + tuple(local.shape))

    def verify(self, draft: torch.Tensor, target: torch.Tensor, tokens: torch.Tensor, r: torch.Tensor,
               noise: torch.Tensor, temperature: float = 1.0,
               draft_temperature: Optional[float] = None) -> Dict[str, torch.Tensor]:
        """draft [B, >=gamma, Vs], target [B, >=gamma+1, Vs], noise [B, Vs]: this rank's slices;
        tokens [B, gamma] (global ids), r [B, gamma]: replicated.  Returns the dict of verify_fused."""
        B, gamma = tokens.shape
        d_temp = temperature if draft_temperature is None else draft_temperature
        if target.shape[-1] != self.v_end - self.v_begin:
            raise ValueError("target slice width does not match this rank's vocabulary slice")
        be = self.backend
        stats = be.row_stats(draft, target, tokens, self.v_begin, temperature, d_temp)
        out, plan = be.accept(self._all_gather(stats), B, gamma, r)
        partial = be.sample_partial(draft, target, plan, noise, self.v_begin, temperature, d_temp, gamma)
        out["token"] = be.finish(self._all_gather(partial), B)
        return out
