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

With ``top_p > 0`` a distributed radix select of every row's nucleus boundary runs between the first gather and
``accept`` (csrc/shard_nucleus.cuh): per level (four for bf16 rows at temperature 1, eight otherwise) a 16-bin
(mass, count) histogram of the slice, an all-gather of the bins (192 B per row) and a reduction to the next key prefix; then the
index cut among equal values (one more 4-byte all-gather) - ``2 + levels + 1`` exchanges per step.  Masses are
fixed-point integers, so the kept set of a row is the same for every world size.

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
        self.PLAN, self.PARTIAL, self.HIST, self.SEL, self.CUT = 48, 24, 192, 32, 16

    def _rows(self, t):
        return (t.data_ptr(), t.stride(0), t.stride(1)) if t is not None and t.numel() else (None, 0, 0)

    def row_stats(self, draft, target, tokens, v_begin, t_temp, d_temp, stats=None, xchg=None):
        B, gamma = tokens.shape
        self.ops._need_cuda(draft, target, tokens)
        if stats is None and xchg is None:
            stats = torch.empty((B * (2 * gamma + 1), 4), dtype=torch.float32, device=target.device)
        dp, dsb, dss = self._rows(draft)
        tp, tsb, tss = self._rows(target)
        with torch.cuda.device(target.device):
            self.check(self.lib.duodec_shard_stats(dp, tp, self.ops._code(target), dsb, dss, tsb, tss, B, gamma,
                                                   target.shape[-1], int(v_begin), tokens.data_ptr(),
                                                   tokens.stride(0) if gamma else 0, float(t_temp), float(d_temp),
                                                   stats.data_ptr() if stats is not None else None, xchg,
                                                   self.ops._stream()), "duodec_shard_stats")
        return stats

    def accept(self, gathered, B, gamma, r, out=None, plan=None, xchg=None, world=None, tokens=None, cuts=None):
        dev = r.device if gathered is None else gathered.device
        if out is None:
            out = {
                "accept_mask": torch.empty((B, gamma), dtype=torch.bool, device=dev),
                "selected_target_p": torch.empty((B, gamma), dtype=torch.float32, device=dev),
                "selected_draft_p": torch.empty((B, gamma), dtype=torch.float32, device=dev),
                "accepted_len": torch.empty((B,), dtype=torch.int64, device=dev),
                "all_accepted": torch.empty((B,), dtype=torch.bool, device=dev),
            }
        if plan is None:
            plan = torch.empty((B, self.PLAN), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            self.check(self.lib.duodec_shard_accept(gathered.data_ptr() if gathered is not None else None,
                                                    world if gathered is None else gathered.shape[0], B, gamma,
                                                    r.data_ptr(),
                                                    r.stride(0) if gamma else 0, out["accept_mask"].data_ptr(),
                                                    out["selected_target_p"].data_ptr(),
                                                    out["selected_draft_p"].data_ptr(), out["accepted_len"].data_ptr(),
                                                    out["all_accepted"].data_ptr(), plan.data_ptr(), xchg,
                                                    tokens.data_ptr() if cuts is not None and gamma else None,
                                                    tokens.stride(0) if cuts is not None and gamma else 0,
                                                    cuts.data_ptr() if cuts is not None else None,
                                                    self.ops._stream()), "duodec_shard_accept")
        return out, plan

    @staticmethod
    def nslice(B, Vs):
        return max(1, min(-(-Vs // 4096), -(-592 // max(B, 1))))  # >= 4 blocks per SM in total

    def sample_partial(self, draft, target, plan, noise, v_begin, t_temp, d_temp, gamma, partial=None, xchg=None,
                       filtered=False):
        B = target.shape[0]
        nslice = self.nslice(B, target.shape[-1])
        if partial is None and xchg is None:
            partial = torch.empty((B, nslice, self.PARTIAL), dtype=torch.uint8, device=target.device)
        dp, dsb, dss = self._rows(draft)
        tp, tsb, tss = self._rows(target)
        with torch.cuda.device(target.device):
            self.check(self.lib.duodec_shard_sample(dp, tp, self.ops._code(target), dsb, dss, tsb, tss, B, gamma,
                                                    target.shape[-1], int(v_begin), float(t_temp), float(d_temp),
                                                    plan.data_ptr(), noise.data_ptr(), noise.stride(0), nslice,
                                                    partial.data_ptr() if partial is not None else None, xchg,
                                                    1 if filtered else 0, self.ops._stream()), "duodec_shard_sample")
        return partial

    # ---- top-p under sharding (csrc/shard_nucleus.cuh)
    def nucleus_levels(self, target, t_temp, d_temp):
        return int(self.lib.duodec_shard_nucleus_levels(self.ops._code(target), float(t_temp), float(d_temp)))

    def nucleus_level(self, draft, target, gamma, v_begin, t_temp, d_temp, top_p, d_top_p, gstats, sel, level, hist=None):
        B = target.shape[0]
        rows = B * (2 * gamma + 1)
        dev = target.device
        if sel is None:
            sel = torch.zeros((rows, self.SEL), dtype=torch.uint8, device=dev)
        if hist is None:
            hist = torch.empty((rows, self.HIST), dtype=torch.uint8, device=dev)
        dp, dsb, dss = self._rows(draft)
        tp, tsb, tss = self._rows(target)
        with torch.cuda.device(dev):
            self.check(self.lib.duodec_shard_nucleus_level(dp, tp, self.ops._code(target), dsb, dss, tsb, tss, B, gamma,
                                                           target.shape[-1], int(v_begin), float(t_temp), float(d_temp),
                                                           float(top_p), float(d_top_p), gstats.data_ptr(),
                                                           gstats.shape[0], sel.data_ptr(), int(level), hist.data_ptr(),
                                                           self.ops._stream()), "duodec_shard_nucleus_level")
        return sel, hist

    def nucleus_reduce(self, ghist, rank, B, gamma, top_p, d_top_p, sel, level, levels, quota=None):
        if quota is None:
            quota = torch.zeros((B * (2 * gamma + 1),), dtype=torch.int32, device=sel.device)
        with torch.cuda.device(sel.device):
            self.check(self.lib.duodec_shard_nucleus_reduce(ghist.data_ptr(), ghist.shape[0], int(rank), B, gamma,
                                                            float(top_p), float(d_top_p), sel.data_ptr(), int(level),
                                                            int(levels), quota.data_ptr(), self.ops._stream()),
                       "duodec_shard_nucleus_reduce")
        return sel, quota

    def nucleus_tie(self, draft, target, gamma, v_begin, t_temp, d_temp, sel, quota, levels, cand=None):
        B = target.shape[0]
        if cand is None:
            cand = torch.empty((B * (2 * gamma + 1),), dtype=torch.int32, device=target.device)
        dp, dsb, dss = self._rows(draft)
        tp, tsb, tss = self._rows(target)
        with torch.cuda.device(target.device):
            self.check(self.lib.duodec_shard_nucleus_tie(dp, tp, self.ops._code(target), dsb, dss, tsb, tss, B, gamma,
                                                         target.shape[-1], int(v_begin), float(t_temp), float(d_temp),
                                                         sel.data_ptr(), quota.data_ptr(), int(levels), cand.data_ptr(),
                                                         self.ops._stream()), "duodec_shard_nucleus_tie")
        return cand

    def nucleus_close(self, sel, gstats, gcand, B, gamma, levels, cuts=None):
        if cuts is None:
            cuts = torch.empty((B * (2 * gamma + 1), self.CUT), dtype=torch.uint8, device=sel.device)
        with torch.cuda.device(sel.device):
            self.check(self.lib.duodec_shard_nucleus_close(sel.data_ptr(), gstats.data_ptr(), gcand.data_ptr(),
                                                           gstats.shape[0], B, gamma, int(levels), cuts.data_ptr(),
                                                           self.ops._stream()), "duodec_shard_nucleus_close")
        return cuts

    def finish(self, gathered, B, token=None, xchg=None, world=None, nslice=None):
        if token is None:
            token = torch.empty((B,), dtype=torch.int64, device=gathered.device)
        with torch.cuda.device(token.device):
            self.check(self.lib.duodec_shard_finish(gathered.data_ptr() if gathered is not None else None,
                                                    world if gathered is None else gathered.shape[0], B,
                                                    nslice if gathered is None else gathered.shape[2],
                                                    token.data_ptr(), xchg,
                                                    self.ops._stream()), "duodec_shard_finish")
        return token


class ShardedVerifier:
    """Verification step over vocabulary slices; call :meth:`verify` with THIS rank's slices."""

    def __init__(self, V: int, rank: Optional[int] = None, world: Optional[int] = None, group=None, backend=None,
                 fused_exchange: bool = False):
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
        # Exchange: NCCL all-gathers captured in the step's graph by default.  `fused_exchange=True` moves both
        # exchanges into the kernels (peer stores + release/acquire flags over symmetric memory); measured on one
        # 8 x B200 box at C5 it is slower than the captured NCCL all-gathers (174 / 213 us vs 147 us per step,
        # profiles/r02_c5_exchange_ab.txt), so it stays opt-in.
        self.collectives = 0
        self.use_graphs = True
        # `fused_exchange="pull"`: results stay in each rank's symmetric buffer, one symmetric-memory barrier per
        # exchange, consumers read the peers' slabs in place over NVLink (no copy, no NCCL call).
        self.fused = bool(fused_exchange) and self.world > 1 and isinstance(backend, _CudaBackend)
        self.pull = self.fused and fused_exchange == "pull"
        self._bufs: dict = {}
        self._graphs: dict = {}
        self._symm: dict = {}
        self.xchg_mode = 0  # 0: every block pushes its result to the peers; 1: the last block copies the slab
        if self.pull:
            self.xchg_mode = 2

    def slice_of(self, full: torch.Tensor) -> torch.Tensor:
        """This rank's columns of a [..., V] tensor (what a sharded LM head would have produced)."""
        return full[..., self.v_begin:self.v_end]

    def _all_gather(self, local: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        if self.world == 1:
            return local.unsqueeze(0)
        local = local.contiguous()
        if out is None:
            out = torch.empty((self.world * local.shape[0],) + tuple(local.shape[1:]), dtype=local.dtype,
                              device=local.device)
        dist.all_gather_into_tensor(out.view((self.world * local.shape[0],) + tuple(local.shape[1:])), local,
                                    group=self.group)  # rank-major concatenation along dim 0
        self.collectives += 1
        return out.view((self.world,) + tuple(local.shape))

    @property
    def exchange_kind(self) -> str:
        if self.world == 1:
            return "none (one rank)"
        if self.fused and self.pull:
            return "pull: symmetric-memory barrier, consumers read the peers' buffers in place (no copy, no NCCL)"
        return ("fused: peer stores + release/acquire flags over symmetric memory, inside the kernels" if self.fused
                else "nccl all_gather_into_tensor (captured in the step's CUDA graph)")

    @property
    def exchanges_per_step(self) -> int:
        return 2 if self.world > 1 else 0

    def _nucleus(self, draft, target, tokens, gstats, t_temp, d_temp, top_p, d_top_p, buf=None):
        """The rows' nucleus filters (x*, i*, kept mass) from the gathered statistics: `levels` rounds of local
        histogram -> all-gather -> reduction, then the index cut among equal values (one more all-gather)."""
        B, gamma = tokens.shape
        be = self.backend
        nb = None
        if buf is not None:
            nb = buf.get("nucleus")
            if nb is None:
                rows, W, dev = B * (2 * gamma + 1), self.world, target.device
                e = lambda shape, dt: torch.empty(shape, dtype=dt, device=dev)  # noqa: E731
                nb = {"sel": torch.zeros((rows, be.SEL), dtype=torch.uint8, device=dev),
                      "hist": e((rows, be.HIST), torch.uint8), "ghist": e((W, rows, be.HIST), torch.uint8),
                      "quota": torch.zeros((rows,), dtype=torch.int32, device=dev), "cand": e((rows,), torch.int32),
                      "gcand": e((W, rows), torch.int32), "cuts": e((rows, be.CUT), torch.uint8)}
                buf["nucleus"] = nb
        g = (lambda k: nb[k]) if nb is not None else (lambda k: None)
        levels = be.nucleus_levels(target, t_temp, d_temp)
        sel, quota = g("sel"), g("quota")
        for level in range(levels):
            sel, hist = be.nucleus_level(draft, target, gamma, self.v_begin, t_temp, d_temp, top_p, d_top_p, gstats, sel,
                                         level, g("hist"))
            sel, quota = be.nucleus_reduce(self._all_gather(hist, g("ghist")), self.rank, B, gamma, top_p, d_top_p,
                                           sel, level, levels, quota)
        cand = be.nucleus_tie(draft, target, gamma, self.v_begin, t_temp, d_temp, sel, quota, levels, g("cand"))
        return be.nucleus_close(sel, gstats, self._all_gather(cand, g("gcand")), B, gamma, levels, g("cuts"))

    def _symmetric(self, B: int, gamma: int, Vs: int, dev: torch.device):
        """Peer-mapped exchange buffer of this shape (torch symmetric memory: CUDA VMM handles exchanged over the
        process group, every rank maps every peer's buffer) and the duodec_shard_xchg that describes it."""
        key = (B, gamma, Vs, dev.index)
        ent = self._symm.get(key)
        if ent is None:
            import torch.distributed._symmetric_memory as symm_mem
            from ._lib import ShardXchg
            rows, W = B * (2 * gamma + 1), self.world
            nslice = self.backend.nslice(B, Vs)
            off_stats = 1024
            off_partial = off_stats + ((W * rows * 16 + 127) // 128) * 128
            nbytes = off_partial + W * B * nslice * 24
            t = symm_mem.empty(nbytes, dtype=torch.uint8, device=dev)
            t.zero_()
            hdl = symm_mem.rendezvous(t, self.group if self.group is not None else dist.group.WORLD)
            torch.cuda.synchronize(dev)
            dist.barrier(group=self.group)  # every header is zero before anybody publishes
            x = ShardXchg(hdl.buffer_ptrs_dev, self.rank, W, off_stats, off_partial, int(self.xchg_mode))
            ent = (t, hdl, x, nslice)
            self._symm[key] = ent
        return ent

    def _buffers(self, B: int, gamma: int, Vs: int, dev: torch.device) -> dict:
        """Every intermediate of a step, allocated once per (B, gamma, slice width): the step itself allocates
        nothing, so it can be captured in a CUDA graph."""
        key = (B, gamma, Vs, dev.index)
        buf = self._bufs.get(key)
        if buf is None:
            rows, W = B * (2 * gamma + 1), self.world
            nslice = getattr(self.backend, "nslice", lambda b, v: 1)(B, Vs)
            e = lambda shape, dt: torch.empty(shape, dtype=dt, device=dev)  # noqa: E731
            buf = {
                "stats": e((rows, 4), torch.float32), "gstats": e((W, rows, 4), torch.float32),
                "plan": e((B, getattr(self.backend, "PLAN", 48)), torch.uint8), "partial": e((B, nslice, 24), torch.uint8),
                "gpartial": e((W, B, nslice, 24), torch.uint8), "token": e((B,), torch.int64),
                "out": {"accept_mask": e((B, gamma), torch.bool), "selected_target_p": e((B, gamma), torch.float32),
                        "selected_draft_p": e((B, gamma), torch.float32), "accepted_len": e((B,), torch.int64),
                        "all_accepted": e((B,), torch.bool)},
            }
            self._bufs[key] = buf
        return buf

    def _step(self, draft, target, tokens, r, noise, temperature, d_temp, buf=None, top_p=0.0, d_top_p=0.0):
        B, gamma = tokens.shape
        be = self.backend
        nucleus = top_p > 0.0 or d_top_p > 0.0
        if buf is None:  # injected test backends: plain allocating calls
            stats = be.row_stats(draft, target, tokens, self.v_begin, temperature, d_temp)
            gstats = self._all_gather(stats)
            if nucleus:
                cuts = self._nucleus(draft, target, tokens, gstats, temperature, d_temp, top_p, d_top_p)
                out, plan = be.accept(gstats, B, gamma, r, tokens=tokens, cuts=cuts)
                partial = be.sample_partial(draft, target, plan, noise, self.v_begin, temperature, d_temp, gamma,
                                            filtered=True)
            else:
                out, plan = be.accept(gstats, B, gamma, r)
                partial = be.sample_partial(draft, target, plan, noise, self.v_begin, temperature, d_temp, gamma)
            out["token"] = be.finish(self._all_gather(partial), B)
            return out
        if nucleus:  # (always over gathered statistics: the fused exchange has no filter)
            stats = be.row_stats(draft, target, tokens, self.v_begin, temperature, d_temp, stats=buf["stats"])
            gstats = self._all_gather(stats, buf["gstats"])
            cuts = self._nucleus(draft, target, tokens, gstats, temperature, d_temp, top_p, d_top_p, buf)
            out, plan = be.accept(gstats, B, gamma, r, out=buf["out"], plan=buf["plan"], tokens=tokens, cuts=cuts)
            partial = be.sample_partial(draft, target, plan, noise, self.v_begin, temperature, d_temp, gamma,
                                        partial=buf["partial"], filtered=True)
            out = dict(out)
            out["token"] = be.finish(self._all_gather(partial, buf["gpartial"]), B, token=buf["token"])
            return out
        if self.fused and self.world > 1:
            import ctypes
            _, hdl, x, nslice = self._symmetric(B, gamma, target.shape[-1], target.device)
            xp = ctypes.byref(x)
            be.row_stats(draft, target, tokens, self.v_begin, temperature, d_temp, xchg=xp)
            if self.pull:
                hdl.barrier(channel=0)  # every rank's statistics are in its buffer (and visible) past this point
            out, plan = be.accept(None, B, gamma, r, out=buf["out"], plan=buf["plan"], xchg=xp, world=self.world)
            be.sample_partial(draft, target, plan, noise, self.v_begin, temperature, d_temp, gamma, xchg=xp)
            if self.pull:
                hdl.barrier(channel=1)
            out = dict(out)
            out["token"] = be.finish(None, B, token=buf["token"], xchg=xp, world=self.world, nslice=nslice)
            return out
        stats = be.row_stats(draft, target, tokens, self.v_begin, temperature, d_temp, stats=buf["stats"])
        out, plan = be.accept(self._all_gather(stats, buf["gstats"]), B, gamma, r, out=buf["out"], plan=buf["plan"])
        partial = be.sample_partial(draft, target, plan, noise, self.v_begin, temperature, d_temp, gamma,
                                    partial=buf["partial"])
        out = dict(out)
        out["token"] = be.finish(self._all_gather(partial, buf["gpartial"]), B, token=buf["token"])
        return out

    def phase_times(self, inputs, temperature: float = 1.0, reps: int = 10) -> Optional[dict]:
        """Measurement hook: microseconds of each kernel / exchange of a step (eager launches, CUDA events)."""
        if not isinstance(self.backend, _CudaBackend) or self.fused:
            return None  # (with the fused exchange the kernels of a step wait on the peers: no stand-alone phases)
        draft, target, tokens, r, noise = inputs
        B, gamma = tokens.shape
        be, buf = self.backend, self._buffers(B, gamma, target.shape[-1], target.device)
        phases = [
            ("stats", lambda: be.row_stats(draft, target, tokens, self.v_begin, temperature, temperature, stats=buf["stats"])),
            ("exchange1", lambda: self._all_gather(buf["stats"], buf["gstats"])),
            ("accept", lambda: be.accept(self._all_gather(buf["stats"], buf["gstats"]) if self.world == 1 else buf["gstats"],
                                         B, gamma, r, out=buf["out"], plan=buf["plan"])),
            ("sample", lambda: be.sample_partial(draft, target, buf["plan"], noise, self.v_begin, temperature,
                                                 temperature, gamma, partial=buf["partial"])),
            ("exchange2", lambda: self._all_gather(buf["partial"], buf["gpartial"])),
            ("finish", lambda: be.finish(buf["gpartial"] if self.world > 1 else buf["partial"].unsqueeze(0), B,
                                         token=buf["token"])),
        ]
        self._step(draft, target, tokens, r, noise, temperature, temperature, buf)
        out = {}
        for name, fn in phases:
            fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                fn()
            e1.record()
            torch.cuda.synchronize()
            out[name] = e0.elapsed_time(e1) * 1e3 / reps
        return out

    def verify(self, draft: torch.Tensor, target: torch.Tensor, tokens: torch.Tensor, r: torch.Tensor,
               noise: torch.Tensor, temperature: float = 1.0, draft_temperature: Optional[float] = None,
               graph: Optional[bool] = None, top_p: float = 0.0,
               draft_top_p: Optional[float] = None) -> Dict[str, torch.Tensor]:
        """draft [B, >=gamma, Vs], target [B, >=gamma+1, Vs], noise [B, Vs]: this rank's slices;
        tokens [B, gamma] (global ids), r [B, gamma]: replicated.  Returns the dict of verify_fused (tensors
        owned by the verifier: valid until the next call with the same shapes).

        With the CUDA backend the four kernels and the two exchanges of a step are captured in ONE CUDA graph
        per set of input buffers (the second call with the same tensors replays it; `graph=False` disables)."""
        B, gamma = tokens.shape
        d_temp = temperature if draft_temperature is None else draft_temperature
        top_p = float(top_p or 0.0)
        d_top_p = top_p if draft_top_p is None else float(draft_top_p or 0.0)
        if target.shape[-1] != self.v_end - self.v_begin:
            raise ValueError("target slice width does not match this rank's vocabulary slice")
        if not isinstance(self.backend, _CudaBackend):
            return self._step(draft, target, tokens, r, noise, temperature, d_temp, None, top_p, d_top_p)
        buf = self._buffers(B, gamma, target.shape[-1], target.device)
        if graph is None:
            graph = self.use_graphs
        if not graph or gamma == 0:
            return self._step(draft, target, tokens, r, noise, temperature, d_temp, buf, top_p, d_top_p)
        key = (draft.data_ptr(), target.data_ptr(), tokens.data_ptr(), r.data_ptr(), noise.data_ptr(),
               tuple(draft.stride()), tuple(target.stride()), B, gamma, target.shape[-1], float(temperature),
               float(d_temp), top_p, d_top_p)
        ent = self._graphs.get(key)
        if ent is None:
            # first call with these buffers: one eager step (also the warm-up NCCL wants), then the capture
            out = self._step(draft, target, tokens, r, noise, temperature, d_temp, buf, top_p, d_top_p)
            torch.cuda.current_stream().synchronize()
            g = torch.cuda.CUDAGraph()
            side = torch.cuda.Stream(device=target.device)
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                with torch.cuda.graph(g, stream=side):
                    out = self._step(draft, target, tokens, r, noise, temperature, d_temp, buf, top_p, d_top_p)
            torch.cuda.current_stream().wait_stream(side)
            if len(self._graphs) >= 16:
                self._graphs.pop(next(iter(self._graphs)))
            # the input tensors are kept alive with the graph that reads them
            self._graphs[key] = (g, out, (draft, target, tokens, r, noise))
            ent = self._graphs[key]
        ent[0].replay()
        return ent[1]


class ShardedTopK:
    """transfer_top_k (``utils.rebuild_topk_uniform_probs``, /root/reference/src/utils.py:751-784) when every rank
    holds a vocabulary slice of the probability rows: a merge of per-shard top-k candidates.

        local     the slice's own top-min(k, Vs) entries           kernel  duodec_topk_rebuild (mode 1)
        compact   (value, global index) lists in index order       kernel  duodec_compact_nonzero
        gather    all ranks' lists: k x 8 B per row and rank        all_gather_into_tensor
        merge     global top-k among the G x k candidates,         kernel  duodec_topk_rebuild (mode 1) on the
                  lowest index on ties                                     gathered values (rank order = index order)
        apply     kept entries keep their value, holes take the     kernel  duodec_shard_topk_apply
                  uniform refill, renormalise

    Every rank ends with its slice of exactly the row the unsharded operator produces."""

    def __init__(self, V: int, rank: Optional[int] = None, world: Optional[int] = None, group=None):
        self.group = group
        self.world = dist.get_world_size(group) if world is None else int(world)
        self.rank = dist.get_rank(group) if rank is None else int(rank)
        self.V = int(V)
        self.v_begin, self.v_end = shard_bounds(self.V, self.rank, self.world)

    # ---- the three local stages (also what the single-device tests drive rank by rank)
    def local_candidates(self, probs: torch.Tensor, k: int, v_begin: Optional[int] = None):
        """probs [rows, Vs] -> (cand_val [rows, kl] f32, cand_idx [rows, kl] i32), kl = min(k, Vs)."""
        from . import _lib, ops
        v_begin = self.v_begin if v_begin is None else v_begin
        rows, Vs = probs.shape
        kl = min(int(k), Vs)
        dense = ops.topk_rebuild(probs, kl, 1) if kl < Vs else probs
        val = torch.empty((rows, kl), dtype=torch.float32, device=probs.device)
        idx = torch.empty((rows, kl), dtype=torch.int32, device=probs.device)
        cnt = torch.empty((rows,), dtype=torch.int32, device=probs.device)
        if rows:
            with torch.cuda.device(probs.device):
                _lib.check(_lib.lib.duodec_compact_nonzero(dense.data_ptr(), ops._code(dense), dense.stride(0), rows, Vs,
                                                           int(v_begin), kl, val.data_ptr(), idx.data_ptr(),
                                                           cnt.data_ptr(), ops._stream()), "duodec_compact_nonzero")
        return val, idx

    def merge(self, vals: torch.Tensor, idxs: torch.Tensor, k: int):
        """Gathered lists [G, rows, kl] (rank-major) -> per-row boundary (tau, istar), refill and normaliser."""
        from . import ops
        G, rows, kl = vals.shape
        v = vals.permute(1, 0, 2).reshape(rows, G * kl).contiguous()
        i = idxs.permute(1, 0, 2).reshape(rows, G * kl)
        picked = ops.topk_rebuild(v.clamp_min(0.0), int(k), 1) if k < G * kl else v.clamp_min(0.0)
        kept = picked > 0
        inf = torch.full_like(v, float("inf"))
        tau = torch.where(kept, v, inf).amin(-1)
        istar = torch.where(kept & (v == tau.unsqueeze(-1)), i, torch.full_like(i, -1)).amax(-1).to(torch.int32)
        S = picked.sum(-1)
        holes = (self.V - kept.sum(-1)).to(torch.float32)
        tail = (1.0 - S).clamp_min(0.0)
        fill = torch.where(holes > 0, tail / holes.clamp_min(1.0), torch.zeros_like(tail))
        total = (S + fill * holes).clamp_min(1e-12)
        return tau.contiguous(), istar.contiguous(), fill.contiguous(), total.contiguous()

    def apply(self, probs: torch.Tensor, boundary, v_begin: Optional[int] = None) -> torch.Tensor:
        from . import _lib, ops
        tau, istar, fill, total = boundary
        v_begin = self.v_begin if v_begin is None else v_begin
        rows, Vs = probs.shape
        out = torch.empty_like(probs)
        if rows:
            with torch.cuda.device(probs.device):
                _lib.check(_lib.lib.duodec_shard_topk_apply(probs.data_ptr(), ops._code(probs), probs.stride(0),
                                                            out.data_ptr(), out.stride(0), rows, Vs, int(v_begin),
                                                            tau.data_ptr(), istar.data_ptr(), fill.data_ptr(),
                                                            total.data_ptr(), ops._stream()), "duodec_shard_topk_apply")
        return out

    def rebuild(self, probs_slice: torch.Tensor, k: Optional[int]) -> torch.Tensor:
        """This rank's slice [..., Vs] of rebuild_topk_uniform_probs(full rows, k)."""
        if k is None or k <= 0 or k >= self.V or probs_slice.numel() == 0:
            return probs_slice
        shape = probs_slice.shape
        p = probs_slice.reshape(-1, shape[-1])
        if p.stride(-1) != 1:
            p = p.contiguous()
        if p.dtype not in (torch.float32, torch.bfloat16):
            p = p.float()
        val, idx = self.local_candidates(p, k)
        if self.world > 1:
            kl = torch.tensor([val.shape[1]], device=p.device)
            kmax = kl.clone()
            dist.all_reduce(kmax, op=dist.ReduceOp.MAX, group=self.group)  # (the last slice may be narrower than k)
            pad = int(kmax) - val.shape[1]
            if pad:
                val = torch.nn.functional.pad(val, (0, pad), value=-1.0)
                idx = torch.nn.functional.pad(idx, (0, pad), value=0x7fffffff)
            gv = torch.empty((self.world,) + tuple(val.shape), dtype=val.dtype, device=p.device)
            gi = torch.empty((self.world,) + tuple(idx.shape), dtype=idx.dtype, device=p.device)
            dist.all_gather_into_tensor(gv.view(-1, val.shape[1]), val.contiguous(), group=self.group)
            dist.all_gather_into_tensor(gi.view(-1, idx.shape[1]), idx.contiguous(), group=self.group)
        else:
            gv, gi = val.unsqueeze(0), idx.unsqueeze(0)
        out = self.apply(p, self.merge(gv, gi, int(k)))
        return out.reshape(shape).to(probs_slice.dtype)
