#!/usr/bin/env python
"""bench.py - the verification hot path at BASELINE.json's config 2.

A step = one fused verification pass (duodec_verify_fused through the C ABI)
over one batch of synthetic logits: B=64 sequences, gamma=4 drafts, V=32000,
fp32, temperature 1.0, top_p 0.9 -> 320 verified positions ("verify tokens").

  value      tokens/s with the inputs resident in HBM, K graph-captured launches
             timed with CUDA events on the launching stream (max over ranks)
  e2e        the same step through HostVerifier with pinned HOST logits: the
             host->device copies and the result read-back are inside the timing
  roofline   logit bytes of one launch / its average duration vs measured HBM copy peak
  cpu_baseline / --impl reference
             the oracle's restatement of the reference torch-CPU path
             (oracle/reference_port.py) on the box's host cores

N>1: every rank verifies its own batch (independent sequences, no collective):
weak scaling, value = all ranks' tokens / max-over-ranks time.
"""

from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

B, GAMMA, V = 64, 4, 32000
TEMP, TOP_K, TOP_P = 1.0, 0, 0.9
SIGMA = 1.0  # draft = target + SIGMA * N(0,1): mid acceptance
NSETS = 4  # rotating input sets: 4 x 90 MB > 126 MB of L2
METRIC = "verify_tokens_per_s"
UNIT = "tokens/s"
WORKLOAD = "C2 synthetic verify: B=64, gamma+1=5, V=32000, fp32 logits, T=1.0, top_p=0.9"


def _config(extra=None):
    c = {
        "workload": WORKLOAD,
        "B": B, "gamma": GAMMA, "V": V, "temperature": TEMP, "top_k": TOP_K, "top_p": TOP_P,
        "draft_sigma": SIGMA,
        "l2": f"{NSETS} rotating input sets ({NSETS * (2 * B * (GAMMA + 1) * V * 4 + B * V * 4) / 1e6:.0f} MB > L2)",
        "sharding": "independent batch per rank, no collective",
    }
    if extra:
        c.update(extra)
    return c


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.tmp = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={index}", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=self.tmp, stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return None
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        self.tmp.flush()
        self.tmp.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.tmp.read().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        self.tmp.close()
        os.unlink(self.tmp.name)
        if not sm:
            return None
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------- inputs
def make_sets(dev, nsets, seed):
    """Seeded synthetic verification inputs on the device."""
    import duodecoding_b200.ops as ops

    g = torch.Generator(device=dev).manual_seed(seed)
    sets = []
    for _ in range(nsets):
        tl = torch.randn(B, GAMMA + 1, V, device=dev, generator=g) * 3
        dl = tl + SIGMA * torch.randn(B, GAMMA + 1, V, device=dev, generator=g)
        q = ops.norm_logits(dl[:, :GAMMA], TEMP, TOP_K, TOP_P)
        e = torch.empty(B * GAMMA, V, device=dev).exponential_(1.0, generator=g)
        toks = ops.sample_rows(q.reshape(B * GAMMA, V), e).reshape(B, GAMMA)
        r = torch.rand(B, GAMMA, device=dev, generator=g)
        noise = torch.empty(B, V, device=dev).exponential_(1.0, generator=g)
        del q, e
        sets.append((dl, tl, toks, r, noise))
    return sets


def cpu_sample(nseq, seed):
    """The same workload for the CPU arms, nseq sequences, generated with the oracle on the host."""
    import oracle

    g = torch.Generator().manual_seed(seed)
    tl = torch.randn(nseq, GAMMA + 1, V, generator=g) * 3
    dl = tl + SIGMA * torch.randn(nseq, GAMMA + 1, V, generator=g)
    q = oracle.norm_logits(dl[:, :GAMMA], TEMP, TOP_K, TOP_P)
    toks = torch.stack(
        [oracle.sample(q[:, j], torch.empty(nseq, V).exponential_(1.0, generator=g))[:, 0] for j in range(GAMMA)], 1)
    r = torch.rand(nseq, GAMMA, generator=g)
    noise = torch.empty(nseq, V).exponential_(1.0, generator=g)
    return dl, tl, toks, r, noise


def time_cpu(nseq, steps, warmup):
    """Oracle port of the reference CPU path, all host threads; returns (tokens/s, s/step, threads)."""
    import oracle

    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    args = cpu_sample(nseq, 1234)
    for _ in range(warmup):
        oracle.fused_verify(*args, TEMP, TOP_K, TOP_P)
    t0 = time.perf_counter()
    for _ in range(steps):
        oracle.fused_verify(*args, TEMP, TOP_K, TOP_P)
    dt = (time.perf_counter() - t0) / max(steps, 1)
    return nseq * (GAMMA + 1) / dt, dt, threads


# ----------------------------------------------------------------------------- reference arm
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    nseq = 16
    v, dt, threads = time_cpu(nseq, args.steps, args.warmup)
    sample = f"{nseq} of the {B} sequences of one C2 batch per step ({nseq * (GAMMA + 1)} positions)"
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": _config({"sample": sample}),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# ----------------------------------------------------------------------------- our arm
def run_ours(args):
    import torch.distributed as dist

    import duodecoding_b200.ops as ops
    from duodecoding_b200.host import HostVerifier

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the verification path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL prints its version banner (NCCL_DEBUG=VERSION/INFO) to stdout: keep stdout to the one JSON line
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    K, W = args.steps, max(args.warmup, 3)
    sets = make_sets(dev, NSETS, 1234 + rank)
    out = None
    for i in range(W):
        out = ops.verify_fused(*sets[i % NSETS], TEMP, TOP_K, TOP_P, out=out)
    torch.cuda.synchronize()
    mean_acc = float(out["accepted_len"].float().mean().item())

    # ---- device-resident: K launches captured in one graph, timed with events on this stream
    stream = torch.cuda.Stream(device=dev)
    with torch.cuda.stream(stream):
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            for i in range(K):
                ops.verify_fused(*sets[i % NSETS], TEMP, TOP_K, TOP_P, out=out)
        graph.replay()  # one untimed replay: graph upload
        stream.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        def load(seconds):  # same launches, untimed: keeps the device under load for the 100 ms clock sampler
            t_end = time.perf_counter() + seconds
            while time.perf_counter() < t_end:
                graph.replay()
                stream.synchronize()

        clocks = ClockSampler(local) if rank == 0 else None
        load(args.clock_seconds)
        barrier()
        e0.record(stream)
        graph.replay()
        e1.record(stream)
        stream.synchronize()
        barrier()
        dt = e0.elapsed_time(e1) * 1e-3
        load(args.clock_seconds)
        clk = clocks.stop() if clocks is not None else None
    # ---- the dominant kernel alone (row scan), same launches, for the roofline line
    ops.set_phase_mask(1)
    try:
        with torch.cuda.stream(stream):
            g1 = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g1, stream=stream):
                for i in range(K):
                    ops.verify_fused(*sets[i % NSETS], TEMP, TOP_K, TOP_P, out=out)
            g1.replay()
            stream.synchronize()
            s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s0.record(stream)
            g1.replay()
            s1.record(stream)
            stream.synchronize()
            scan_us = s0.elapsed_time(s1) * 1e3 / K
    finally:
        ops.set_phase_mask(3)
    dt = max_over_ranks(dt)
    tokens = world * B * (GAMMA + 1) * K
    value = tokens / dt
    us_launch = dt / K * 1e6

    # ---- end to end: pinned host logits -> HostVerifier -> pinned host results
    hv = HostVerifier(B, GAMMA, V, torch.float32, dev, nsplit=4)
    hsets = []
    for s in sets[:2]:
        hsets.append(tuple(t.cpu().pin_memory() for t in s[:4]))
    Ke = max(3, min(K, 50))
    for i in range(3):
        hv.verify(*hsets[i % 2], None, TEMP, TOP_K, TOP_P)
    barrier()
    t0 = time.perf_counter()
    chk = 0
    for i in range(Ke):
        res = hv.verify(*hsets[i % 2], None, TEMP, TOP_K, TOP_P)
        chk += int(res["accepted_len"][0])  # the Python-int read the reference's callers make
    torch.cuda.synchronize()
    dte = time.perf_counter() - t0
    barrier()
    dte = max_over_ranks(dte)
    e2e_value = world * B * (GAMMA + 1) * Ke / dte

    if world > 1:
        dist.barrier()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    # ---- roofline of the one kernel a step launches
    alg_bytes = 2 * B * (GAMMA + 1) * V * 4  # every draft/target logit once (81.92 MB)
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    else:
        peak, peak_src = 6650.0, "B200_PROFILING.md fallback (of fallback)"
    achieved = alg_bytes / (scan_us * 1e-6) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        try:
            traffic = json.load(open(tp)).get("scan_kernel_bytes_per_launch")
        except (ValueError, OSError):
            traffic = None
    roofline = {
        "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
        "traffic": traffic, "kernel": "pair::row_pair_kernel<float>", "alg_bytes_per_launch": alg_bytes,
        "us_per_launch": scan_us, "peak_source": peak_src, "frac_of_8TBps_nominal": achieved / 8000.0,
        "step_us": us_launch, "step_kernels": ["pair::row_pair_kernel<float>", "seq_sample_tma_kernel<float>"],
        "step_frac": alg_bytes / (us_launch * 1e-6) / 1e9 / peak,
    }

    # ---- CPU baseline: oracle port of the reference path on the host cores (bounded sample)
    cpu = None
    if world == 1 and not args.no_cpu:
        nseq, csteps = 16, 40
        v, cdt, threads = time_cpu(nseq, csteps, 2)
        cpu = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": f"{csteps} steps x {nseq} of the {B} sequences of a C2 batch ({cdt * 1e3:.1f} ms/step)"}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": dt / K * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": _config({"mean_accepted_len": mean_acc, "timing": "CUDA events around one graph of K launches"}),
        "clocks": clk,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": hv.h2d_bytes,
                "d2h_bytes_per_step": hv.d2h_bytes, "steps": Ke, "ms_per_step": dte / Ke * 1e3,
                "api": "duodecoding_b200.host.HostVerifier.verify (pinned host logits, 4-way copy/verify pipeline)"},
        "gpu_launches": 2 * K,
        "roofline": roofline,
        "cpu_baseline": cpu,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


# ----------------------------------------------------------------------------- C5: vocabulary-sharded verify
def run_c5(args):
    """BASELINE.json config 5: B=256, gamma=6, V=151936, bf16, no filter, vocabulary sharded over the ranks
    (strong scaling: the total work is fixed, every rank holds V/N columns of every row)."""
    import torch.distributed as dist

    from duodecoding_b200.sharded import ShardedVerifier

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the verification path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL prints its version banner (NCCL_DEBUG=VERSION/INFO) to stdout: keep stdout to the one JSON line
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)
    Bc, Gc, Vc = 256, 6, 151936
    sv = ShardedVerifier(Vc, rank=rank, world=world)
    Vs = sv.v_end - sv.v_begin
    g = torch.Generator(device=dev).manual_seed(4321 + rank)
    gt = torch.Generator(device=dev).manual_seed(99)  # replicated tensors: same seed on every rank
    sets = []
    for _ in range(2 if world == 1 else 4):
        tl = (torch.randn(Bc, Gc + 1, Vs, device=dev, generator=g) * 3).to(torch.bfloat16)
        dl = (tl[:, :Gc].float() + torch.randn(Bc, Gc, Vs, device=dev, generator=g)).to(torch.bfloat16)
        toks = torch.randint(0, Vc, (Bc, Gc), device=dev, generator=gt)
        r = torch.rand(Bc, Gc, device=dev, generator=gt) * 0.05  # low thresholds: a mix of accepted lengths
        noise = torch.empty(Bc, Vs, device=dev).exponential_(1.0, generator=g)
        sets.append((dl, tl, toks, r, noise))
    K, W = args.steps, max(args.warmup, 3)
    for i in range(W):
        out = sv.verify(*sets[i % len(sets)], 1.0)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    clocks = ClockSampler(local) if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(K):
        out = sv.verify(*sets[i % len(sets)], 1.0)
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    dt = e0.elapsed_time(e1) * 1e-3
    if world > 1:
        t = torch.tensor([dt], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    # end to end through the public API: the replicated small inputs come from pinned host memory, the
    # results go back to the host every step (the logit slices are produced on the device by the LM head)
    h_toks, h_r = sets[0][2].cpu().pin_memory(), sets[0][3].cpu().pin_memory()
    t0 = time.perf_counter()
    Ke = max(3, min(K, 20))
    for i in range(Ke):
        s = sets[i % len(sets)]
        res = sv.verify(s[0], s[1], h_toks.to(dev, non_blocking=True), h_r.to(dev, non_blocking=True), s[4], 1.0)
        host = {k: v.cpu() for k, v in res.items()}
        _ = int(host["accepted_len"][0])
    dte = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([dte], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dte = float(t.item())
    clk = clocks.stop() if clocks is not None else None
    if rank == 0:
        alg = 2 * Bc * (Gc + 1) * Vc * 2
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        peak = float(json.load(open(peaks_path))["hbm_gbs"]) if os.path.exists(peaks_path) else 6650.0
        ach = alg / world / (dt / K) / 1e9
        line = {
            "metric": METRIC, "value": Bc * (Gc + 1) * K / dt, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": dt / K * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "C5 vocabulary-sharded verify: B=256, gamma+1=7, V=151936, bf16 logits, T=1.0, "
                                   "no filter", "B": Bc, "gamma": Gc, "V": Vc, "sharding": f"vocab/{world}",
                       "collectives_per_step": 2 if world > 1 else 0, "mean_accepted_len":
                           float(out["accepted_len"].float().mean().item()),
                       "l2": "per-rank inputs larger than L2 or rotated", "timing": "CUDA events around K steps"},
            "clocks": clk, "gpu_launches": 4 * K,
            "e2e": {"value": Bc * (Gc + 1) * Ke / dte, "unit": UNIT, "h2d_bytes_per_step": int(h_toks.numel() * 8 +
                    h_r.numel() * 4), "d2h_bytes_per_step": int(sum(v.numel() * v.element_size() for v in host.values())),
                    "api": "duodecoding_b200.sharded.ShardedVerifier.verify"},
            "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                         "traffic": None, "kernel": "shard_stats_kernel<bf16> (+accept, sample, finish, 2 all-gathers)",
                         "alg_bytes_per_launch": alg // world, "us_per_launch": dt / K * 1e6,
                         "note": "per-GPU bytes of one step / whole step time"},
            "cpu_baseline": None,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=["c2", "c5"], help="c2: BASELINE config 2 (default, batch per rank); c5: vocabulary-sharded config 5")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--clock-seconds", type=float, default=0.6,
                    help="untimed replays of the same graph before and after the timed one, for the clock sampler")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if args.workload == "c5":
        return run_c5(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
