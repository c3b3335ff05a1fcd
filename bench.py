#!/usr/bin/env python
"""bench.py - the verification hot path at BASELINE.json's configs.

Default workload (`--workload c2`, BASELINE config 2, the one `metric` is quoted on): a step = one
fused verification pass (duodec_verify_fused through the C ABI) over one batch of synthetic logits:
B=64 sequences, gamma=4 drafts, V=32000, fp32, temperature 1.0, top_p 0.9 -> 320 verified positions
("verify tokens").

  value        tokens/s with the inputs (logits, r, pre-drawn Exp(1) noise) resident in HBM, K launches
               captured in one graph, CUDA events on the launching stream (max over ranks); the step's
               result is compared with the CPU oracle on rank 0 BEFORE timing (`checked`)
  config.step_with_rng_us / isolated_step_us
               the same step with the r / noise draws inside the captured step, and replayed one launch
               at a time (no overlap of consecutive steps' kernels)
  api          one step through the reference-named operator on device-resident KVCacheModel histories:
               decoding_ops.resolve_stage_verification (draws, fused kernels, the Python-int read-back,
               rollback) - what a caller of /root/reference/src/decoding_ops.py:353-397 gets
  e2e          the same step through HostVerifier with pinned HOST logits: host->device copies and the
               result read-back inside the timing
  roofline     logit bytes of one launch / the scan kernel's average duration vs measured HBM copy peak
  cpu_baseline / --impl reference
               the oracle's restatement of the reference torch-CPU path (oracle/reference_port.py): one
               batched norm_logits per model, accept test, per-row resample, FULL batch, all host threads

Other workloads: `c3` (B=32, bf16, `--temperature 0|1`), `c4` (two-stage tridecoding verification with
transfer_top_k=1024 through the mirrored operators), `c5` (B=256, gamma=6, V=151936 bf16, vocabulary
sharded over the ranks).  N>1 with the default workload: every rank verifies its own batch (independent
sequences, no collective: weak scaling) and the line carries a `c5` sub-record measured at that N.
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

_STDOUT = None


def emit(line: dict) -> None:
    text = json.dumps(line) + "\n"
    if _STDOUT is None:
        sys.stdout.write(text)
        sys.stdout.flush()
    else:
        os.write(_STDOUT, text.encode())


METRIC = "verify_tokens_per_s"
UNIT = "tokens/s"
NSETS = 4  # rotating input sets: > 126 MB of L2 in total for c2
TEMP, TOP_K, TOP_P = 1.0, 0, 0.9  # the default workload's warp (tools/ read these)


class Workload:
    def __init__(self, name, B, gamma, V, dtype, T, top_k, top_p, sigma, label):
        self.name, self.B, self.gamma, self.V, self.dtype = name, B, gamma, V, dtype
        self.T, self.top_k, self.top_p, self.sigma, self.label = T, top_k, top_p, sigma, label

    @property
    def elt(self):
        return 4 if self.dtype == torch.float32 else 2

    @property
    def alg_bytes(self):  # every draft / target logit once
        return 2 * self.B * (self.gamma + 1) * self.V * self.elt

    @property
    def set_bytes(self):
        return self.alg_bytes + self.B * self.V * 4

    def config(self, extra=None):
        c = {
            "workload": self.label, "B": self.B, "gamma": self.gamma, "V": self.V, "temperature": self.T,
            "top_k": self.top_k, "top_p": self.top_p, "draft_sigma": self.sigma,
            "l2": f"{NSETS} rotating input sets ({NSETS * self.set_bytes / 1e6:.0f} MB > L2)",
            "sharding": "independent batch per rank, no collective",
        }
        if extra:
            c.update(extra)
        return c


def workload(args) -> Workload:
    if args.workload == "c3":
        T = 1.0 if args.temperature is None else float(args.temperature)
        return Workload("c3", 32, 4, 32000, torch.bfloat16, T, 0, 0.0, 0.7,
                        f"C3 shapes, synthetic verify: B=32, gamma+1=5, V=32000, bf16 logits, T={T}, no filter")
    return Workload("c2", 64, 4, 32000, torch.float32, 1.0, 0, 0.9, 1.0,
                    "C2 synthetic verify: B=64, gamma+1=5, V=32000, fp32 logits, T=1.0, top_p=0.9")


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


def hbm_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        return float(json.load(open(path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    return 6650.0, "B200_PROFILING.md fallback (of fallback)"


def finish_ranks(dist):
    """End of a multi-rank run: one last barrier, then leave without tearing the communicators down (a teardown
    that hangs would hold eight GPUs until the caller's timeout)."""
    try:
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
    finally:
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


def dist_env():
    return int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))


def init_dist(dev, world):
    import torch.distributed as dist

    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL prints its version banner (NCCL_DEBUG=VERSION/INFO) to stdout: keep stdout to the one JSON line
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)
    return dist


# ----------------------------------------------------------------------------- inputs
def make_sets(dev, nsets, seed, wl: Workload = None):
    """Seeded synthetic verification inputs on the device: target = 3 N(0,1), draft = target +
    sigma N(0,1), drafted tokens sampled from the warped draft rows, r ~ U[0,1), Exp(1) noise."""
    import duodecoding_b200.ops as ops

    if wl is None:
        wl = Workload("c2", 64, 4, 32000, torch.float32, 1.0, 0, 0.9, 1.0, "")
    B, G, V = wl.B, wl.gamma, wl.V
    g = torch.Generator(device=dev).manual_seed(seed)
    sets = []
    for _ in range(nsets):
        tl = torch.randn(B, G + 1, V, device=dev, generator=g) * 3
        dl = tl + wl.sigma * torch.randn(B, G + 1, V, device=dev, generator=g)
        tl, dl = tl.to(wl.dtype), dl.to(wl.dtype)
        q = ops.norm_logits(dl[:, :G], wl.T, wl.top_k, wl.top_p)
        e = torch.empty(B * G, V, device=dev).exponential_(1.0, generator=g)
        toks = ops.sample_rows(q.reshape(B * G, V), e).reshape(B, G)
        r = torch.rand(B, G, device=dev, generator=g)
        noise = torch.empty(B, V, device=dev).exponential_(1.0, generator=g)
        del q, e
        sets.append((dl, tl, toks, r, noise))
    return sets


def check_against_oracle(wl, s, got):
    """One step's discrete outputs against the CPU oracle (rank 0, before anything is timed)."""
    import oracle

    dl, tl, toks, r, noise = (t.cpu() for t in s)
    want = oracle.fused_verify(dl[:, :wl.gamma].float(), tl.float(), toks, r, noise, wl.T, wl.top_k, wl.top_p)
    ok = all(torch.equal(got[k].cpu(), want[k]) for k in ("accepted_len", "token", "accept_mask"))
    if not ok:
        raise SystemExit(f"bench.py: the {wl.name} step does not match the CPU oracle - refusing to time it")
    return True


def time_cpu(wl, steps, warmup):
    """Oracle port of the reference CPU path on the FULL batch, all host threads: (tokens/s, s/step, threads)."""
    import oracle

    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    g = torch.Generator().manual_seed(1234)
    B, G, V = wl.B, wl.gamma, wl.V
    tl = (torch.randn(B, G + 1, V, generator=g) * 3).to(wl.dtype).float()
    dl = (tl[:, :G] + wl.sigma * torch.randn(B, G, V, generator=g)).to(wl.dtype).float()
    q = oracle.norm_logits(dl, wl.T, wl.top_k, wl.top_p)
    toks = torch.stack([oracle.sample(q[:, j], torch.empty(B, V).exponential_(1.0, generator=g))[:, 0]
                        for j in range(G)], 1)
    r = torch.rand(B, G, generator=g)
    noise = torch.empty(B, V).exponential_(1.0, generator=g)
    for _ in range(warmup):
        oracle.fused_verify(dl, tl, toks, r, noise, wl.T, wl.top_k, wl.top_p)
    t0 = time.perf_counter()
    for _ in range(steps):
        oracle.fused_verify(dl, tl, toks, r, noise, wl.T, wl.top_k, wl.top_p)
    dt = (time.perf_counter() - t0) / max(steps, 1)
    return B * (G + 1) / dt, dt, threads


# ----------------------------------------------------------------------------- reference arm
def run_reference(args):
    """The reference's own CPU implementation of the path (oracle port: the Python reference cannot travel to
    the box) on this arm's config, all host threads; each step is the full batch."""
    world, rank, _ = dist_env()
    if rank != 0:
        return 0
    wl = workload(args)
    steps = max(1, min(args.steps, 30))  # ~0.3 s per full C2 batch: a bounded sample of the K steps asked for
    v, dt, threads = time_cpu(wl, steps, min(max(args.warmup, 1), 3))
    sample = f"{steps} steps of the full batch ({wl.B} sequences, {wl.B * (wl.gamma + 1)} positions, {dt * 1e3:.1f} ms/step)"
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32" if wl.dtype == torch.float32 else "bf16", "data": "synthetic",
        "config": wl.config({"sample": sample}),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# ----------------------------------------------------------------------------- the reference-named operator
class _NoModel:
    """What KVCacheModel's constructor needs from a model; the bench loads logit rows by hand."""

    class _Emb:
        def __init__(self, V, dev):
            self.weight = torch.zeros(V, 1, device=dev)

    class _Cfg:
        def __init__(self, V):
            self.vocab_size = V

    def __init__(self, V, dev):
        self.config, self._emb = self._Cfg(V), self._Emb(V, dev)

    def get_input_embeddings(self):
        return self._emb

    def parameters(self):
        return iter([self._emb.weight])


class _Len:
    """Transformer-cache stand-in: a length with crop / get_seq_length (src/model_gpu.py:23-26)."""

    def __init__(self, n):
        self.n = n

    def crop(self, end):
        self.n = min(self.n, end)

    def get_seq_length(self, layer_idx=0):
        return self.n


def api_leg(wl, s, dev, steps):
    """decoding_ops.resolve_stage_verification on KVCacheModel caches whose histories hold this step's logit
    rows (device resident).  Per step: r and Exp(1) draws on the device, the fused kernels, the Python-int
    read-back of the accepted count, both rollbacks - the whole reference-named call, timed on the host clock
    around `steps` calls with a final synchronize."""
    from duodecoding_b200 import decoding_ops, model_gpu

    dl, tl, toks, _, _ = s
    B, G, V, P = wl.B, wl.gamma, wl.V, 8
    dh = torch.cat((dl[:, :1].expand(B, P - 1, V), dl[:, :G]), 1).contiguous()
    th = torch.cat((tl[:, :1].expand(B, P - 1, V), tl[:, :G + 1]), 1).contiguous()
    x = torch.cat((torch.zeros(B, P, dtype=torch.int64, device=dev), toks), 1)
    dc = model_gpu.KVCacheModel(_NoModel(V, dev), wl.T, wl.top_k, wl.top_p)
    tc = model_gpu.KVCacheModel(_NoModel(V, dev), wl.T, wl.top_k, wl.top_p)
    dc._store_rows(dh, 0)
    tc._store_rows(th, 0)
    dc._past_key_values, tc._past_key_values = _Len(P - 1 + G), _Len(P + G)
    del dh, th

    def step():
        dc._current_seq_len, dc._past_key_values.n = P - 1 + G, P - 1 + G
        tc._current_seq_len, tc._past_key_values.n = P + G, P + G
        return decoding_ops.resolve_stage_verification(dc, tc, x, P, G, output_device=dev)

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    acc = 0
    for _ in range(steps):
        acc += step()[0]
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / steps
    assert dc._prob_buffer is None and tc._prob_buffer is None, "the fused route must not materialise probabilities"
    return {"name": "duodecoding_b200.decoding_ops.resolve_stage_verification (KVCacheModel logit histories, fused route)",
            "reference": "/root/reference/src/decoding_ops.py:353-397", "us_per_step": dt * 1e6,
            "value": B * (G + 1) / dt, "unit": UNIT, "steps": steps, "mean_accepted_count_seq0": acc / steps,
            "includes": "torch.rand + exponential_ draws, fused kernels, .item() read-back, rollbacks",
            "alg_GBps": wl.alg_bytes / dt / 1e9}


def graph_time(stream, build, K, load_seconds=0.0, barrier=None):
    """Capture `build()` (which launches K steps) in a graph on `stream`, replay once untimed, then time one
    replay with CUDA events on that stream; returns seconds."""
    with torch.cuda.stream(stream):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=stream):
            build()
        g.replay()
        stream.synchronize()
        t_end = time.perf_counter() + load_seconds
        while time.perf_counter() < t_end:
            g.replay()
            stream.synchronize()
        if barrier is not None:
            barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        g.replay()
        e1.record(stream)
        stream.synchronize()
        if barrier is not None:
            barrier()
    return e0.elapsed_time(e1) * 1e-3, g


# ----------------------------------------------------------------------------- our arm (c2 / c3)
def run_ours(args):
    import duodecoding_b200.ops as ops
    from duodecoding_b200.host import HostVerifier

    world, rank, local = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the verification path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = init_dist(dev, world)
    wl = workload(args)

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
    warp = (wl.T, wl.top_k, wl.top_p)
    sets = make_sets(dev, NSETS, 1234 + rank, wl)
    out = None
    for i in range(W):
        out = ops.verify_fused(*sets[i % NSETS], *warp, out=out)
    torch.cuda.synchronize()
    checked = False
    if rank == 0 and not args.no_check:
        got = ops.verify_fused(*sets[0], *warp)
        torch.cuda.synchronize()
        checked = check_against_oracle(wl, sets[0], got)
    out = ops.verify_fused(*sets[(W - 1) % NSETS], *warp, out=out)
    torch.cuda.synchronize()
    mean_acc = float(out["accepted_len"].float().mean().item())

    # ---- device-resident: K launches captured in one graph, timed with events on this stream
    stream = torch.cuda.Stream(device=dev)

    def steps_plain():
        for i in range(K):
            ops.verify_fused(*sets[i % NSETS], *warp, out=out)

    clocks = ClockSampler(local) if rank == 0 else None
    dt, graph = graph_time(stream, steps_plain, K, args.clock_seconds, barrier)
    with torch.cuda.stream(stream):
        t_end = time.perf_counter() + args.clock_seconds
        while time.perf_counter() < t_end:
            graph.replay()
            stream.synchronize()
    clk = clocks.stop() if clocks is not None else None
    dt = max_over_ranks(dt)
    tokens = world * wl.B * (wl.gamma + 1) * K
    value = tokens / dt
    us_launch = dt / K * 1e6

    # ---- the dominant kernel alone (row scan / arg-max), same launches, for the roofline line
    greedy = wl.T == 0.0
    ops.set_phase_mask(1)
    try:
        sdt, _ = graph_time(stream, steps_plain, K)
        scan_us = sdt / K * 1e6
    finally:
        ops.set_phase_mask(3)
    if greedy:
        scan_us = us_launch  # the greedy path has no phase split: memset + arg-max + close

    # ---- the same step with its randomness drawn inside the captured step, and one launch per replay
    def steps_rng():
        for i in range(K):
            s = sets[i % NSETS]
            s[3].uniform_()
            s[4].exponential_(1.0)
            ops.verify_fused(*s, *warp, out=out)

    rdt, _ = graph_time(stream, steps_rng, K)
    one_dt, one = graph_time(stream, lambda: ops.verify_fused(*sets[0], *warp, out=out), 1)
    with torch.cuda.stream(stream):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(50):
            one.replay()
        e1.record(stream)
        stream.synchronize()
    isolated_us = e0.elapsed_time(e1) * 1e3 / 50

    # ---- the reference-named operator on device-resident histories
    api = api_leg(wl, sets[0], dev, max(10, min(K, 100))) if rank == 0 else None

    # ---- end to end: pinned host logits -> HostVerifier -> pinned host results
    hv = HostVerifier(wl.B, wl.gamma, wl.V, wl.dtype, dev)
    hsets = []
    for s in sets[:2]:  # two pinned input sets (the producer's double buffer), filled before the timed region
        hs = hv.host_set()
        hs["draft"].copy_(s[0][:, :wl.gamma])
        hs["target"].copy_(s[1][:, :wl.gamma + 1])
        hs["tokens"].copy_(s[2])
        hs["r"].copy_(s[3])
        hsets.append(hs)
    torch.cuda.synchronize()
    Ke = max(3, min(K, 50))
    for i in range(3):
        hv.verify_set(hsets[i % 2], False, *warp)
    barrier()
    t0 = time.perf_counter()
    chk = 0
    for i in range(Ke):
        res = hv.verify_set(hsets[i % 2], False, *warp)
        chk += int(res["accepted_len"][0])  # the Python-int read the reference's callers make
    torch.cuda.synchronize()
    dte = time.perf_counter() - t0
    barrier()
    dte = max_over_ranks(dte)
    e2e_value = world * wl.B * (wl.gamma + 1) * Ke / dte
    h2d_bytes, d2h_bytes = hv.h2d_bytes, hv.d2h_bytes
    del hv, hsets

    # ---- N > 1: the vocabulary-sharded config 5 at this N rides along as a sub-record
    c5 = None
    if world > 1 and not args.no_c5:
        del sets
        torch.cuda.empty_cache()
        c5 = c5_measure(dev, world, rank, dist, max(10, min(K, 50)), 3)
    if world > 1:
        dist.barrier()
    if rank != 0:
        finish_ranks(dist)
        return 0

    # ---- roofline of the dominant kernel
    peak, peak_src = hbm_peak()
    achieved = wl.alg_bytes / (scan_us * 1e-6) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp) and wl.name == "c2":
        try:
            traffic = json.load(open(tp)).get("scan_kernel_bytes_per_launch")
        except (ValueError, OSError):
            traffic = None
    tname = "float" if wl.dtype == torch.float32 else "__nv_bfloat16"
    kernels = ([f"greedy::greedy_argmax_kernel<{tname}>", "greedy::greedy_close_kernel"] if greedy else
               [f"pair::row_pair_kernel<{tname}>", f"seq_sample_tma_kernel<{tname}>"])
    roofline = {
        "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
        "traffic": traffic, "kernel": kernels[0], "alg_bytes_per_launch": wl.alg_bytes,
        "us_per_launch": scan_us, "peak_source": peak_src, "frac_of_8TBps_nominal": achieved / 8000.0,
        "step_us": us_launch, "step_kernels": kernels,
        "step_frac": wl.alg_bytes / (us_launch * 1e-6) / 1e9 / peak,
    }

    # ---- CPU baseline: oracle port of the reference path on the host cores, full batch
    cpu = None
    if world == 1 and not args.no_cpu:
        csteps = 30 if wl.name == "c2" else 60
        v, cdt, threads = time_cpu(wl, csteps, 2)
        cpu = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": f"{csteps} steps of the full batch ({wl.B} sequences, {cdt * 1e3:.1f} ms/step)"}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": dt / K * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32" if wl.dtype == torch.float32 else "bf16", "data": "synthetic",
        "config": wl.config({"mean_accepted_len": mean_acc, "checked": checked,
                             "timing": "CUDA events around one graph of K launches; r and noise pre-drawn",
                             "step_with_rng_us": rdt / K * 1e6, "isolated_step_us": isolated_us}),
        "clocks": clk,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes, "steps": Ke,
                "ms_per_step": dte / Ke * 1e3,
                "api": "duodecoding_b200.host.HostVerifier.verify_set (pinned host arena: one H2D copy, fused verify, one D2H copy, one sync)"},
        "gpu_launches": (3 if greedy else 2) * K,
        "api": api,
        "roofline": roofline,
        "cpu_baseline": cpu,
    }
    if c5 is not None:
        line["c5"] = c5
    emit(line)
    if world > 1:
        finish_ranks(dist)
    return 0


# ----------------------------------------------------------------------------- C5: vocabulary-sharded verify
def c5_measure(dev, world, rank, dist, K, W, top_p=0.0):
    """BASELINE.json config 5: B=256, gamma=6, V=151936, bf16, no filter, vocabulary sharded over the ranks
    (strong scaling: total work fixed, every rank holds V/N columns of every row).  Every rank builds the same
    full tensors from a shared seed and keeps its columns; rank 0 also verifies the full tensors on one GPU and
    requires the sharded result to be identical (`checked`).  Returns the sub-record (None off rank 0).
    `top_p` > 0 (bench.py --workload c5 --top-p 0.9): the same step with the nucleus filter on both models - the
    distributed radix select of duodecoding_b200/csrc/shard_nucleus.cuh (four 4-bit levels for bf16 rows at T = 1)."""
    import duodecoding_b200.ops as ops
    from duodecoding_b200.sharded import ShardedVerifier

    Bc, Gc, Vc = 256, 6, 151936
    sv = ShardedVerifier(Vc, rank=rank, world=world, fused_exchange=os.environ.get("DUODEC_C5_EXCHANGE") or False)
    g = torch.Generator(device=dev).manual_seed(4321)  # the same full tensors on every rank
    sets, full0 = [], None
    for si in range(2):
        tl = (torch.randn(Bc, Gc + 1, Vc, device=dev, generator=g) * 3).to(torch.bfloat16)
        dl = (tl[:, :Gc].float() + torch.randn(Bc, Gc, Vc, device=dev, generator=g)).to(torch.bfloat16)
        q = ops.norm_logits(dl, 1.0, 0, top_p)
        e = torch.empty(Bc * Gc, Vc, device=dev).exponential_(1.0, generator=g)
        toks = ops.sample_rows(q.reshape(Bc * Gc, Vc), e).reshape(Bc, Gc)  # drafted from q: both branches occur
        del q, e
        r = torch.rand(Bc, Gc, device=dev, generator=g)
        noise = torch.empty(Bc, Vc, device=dev).exponential_(1.0, generator=g)
        if si == 0 and rank == 0:
            full0 = ops.verify_fused(dl, tl, toks, r, noise, 1.0, 0, top_p)
            full0 = {k: v.clone() for k, v in full0.items()}
        sets.append((sv.slice_of(dl).contiguous(), sv.slice_of(tl).contiguous(), toks, r,
                     sv.slice_of(noise).contiguous()))
        del tl, dl, noise
    torch.cuda.empty_cache()
    out = sv.verify(*sets[0], 1.0, top_p=top_p)
    torch.cuda.synchronize()
    checked = False
    if rank == 0:
        checked = all(torch.equal(out[k], full0[k]) for k in ("accepted_len", "token", "accept_mask"))
        if not checked:
            raise SystemExit("bench.py: the sharded C5 step differs from the single-GPU result - refusing to time it")
    for i in range(W):
        out = sv.verify(*sets[i % 2], 1.0, top_p=top_p)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(K):
        out = sv.verify(*sets[i % 2], 1.0, top_p=top_p)
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    dt = e0.elapsed_time(e1) * 1e-3
    if world > 1:
        t = torch.tensor([dt], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    mean_acc = float(out["accepted_len"].float().mean().item())
    timings = sv.phase_times(sets[0], 1.0) if hasattr(sv, "phase_times") and not top_p else None
    nuc_x = 5 if top_p else 0  # bf16 rows at T = 1: four levels + the index cut
    if rank != 0:
        return None
    alg = 2 * Bc * (Gc + 1) * Vc * 2
    peak, peak_src = hbm_peak()
    ach = alg / world / (dt / K) / 1e9
    return {
        "workload": "C5 vocabulary-sharded verify: B=256, gamma+1=7, V=151936, bf16 logits, T=1.0, "
                    + (f"top_p={top_p} on both models (distributed radix select)" if top_p else "no filter"),
        "top_p": top_p, "launches_per_step": 4 + (10 if top_p else 0),
        "nranks": world, "scaling": "strong", "ms_per_step": dt / K * 1e3, "value": Bc * (Gc + 1) * K / dt,
        "unit": UNIT, "steps": K, "checked": checked, "mean_accepted_len": mean_acc,
        "exchanges_per_step": (sv.exchanges_per_step + nuc_x) if world > 1 else 0,
        "exchange": getattr(sv, "exchange_kind", "nccl all_gather_into_tensor"),
        "phase_us": timings,
        "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None,
                     "kernel": "shard_stats_kernel<__nv_bfloat16> (+ accept, sample, finish, exchanges)",
                     "alg_bytes_per_launch": alg // world, "us_per_launch": dt / K * 1e6, "peak_source": peak_src,
                     "note": "per-GPU logit bytes of one step / whole step time, max over ranks"},
    }


def run_c5(args):
    world, rank, local = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the verification path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = init_dist(dev, world)
    clocks = ClockSampler(local) if rank == 0 else None
    rec = c5_measure(dev, world, rank, dist, args.steps, max(args.warmup, 3), float(args.top_p or 0.0))
    clk = clocks.stop() if clocks is not None else None
    if rank == 0:
        line = {
            "metric": METRIC, "value": rec["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": rec["ms_per_step"], "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": rec["workload"], "B": 256, "gamma": 6, "V": 151936, "sharding": f"vocab/{world}",
                       "exchanges_per_step": rec["exchanges_per_step"], "exchange": rec["exchange"],
                       "mean_accepted_len": rec["mean_accepted_len"], "checked": rec["checked"],
                       "phase_us": rec["phase_us"], "l2": "per-rank inputs larger than L2 or rotated",
                       "timing": "CUDA events around K steps"},
            "clocks": clk, "gpu_launches": rec["launches_per_step"] * args.steps,
            "e2e": {"value": rec["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                    "note": "the logit slices are produced on the device by the sharded LM head: no host leg"},
            "roofline": rec["roofline"], "cpu_baseline": None,
        }
        emit(line)
    if world > 1:
        finish_ranks(dist)
    return 0


# ----------------------------------------------------------------------------- C4: two-stage tridecoding verify
C4 = {"B": 8, "V": 32000, "gamma2": 4, "gamma1": 9, "k": 1024, "P": 8, "T": 1.0}


def _c4_logits(dev, seed):
    g = torch.Generator(device=dev).manual_seed(seed)
    B, V, P, g2, g1 = C4["B"], C4["V"], C4["P"], C4["gamma2"], C4["gamma1"]
    L = P + g1 + 1
    target = (torch.randn(B, L, V, device=dev, generator=g) * 3).to(torch.bfloat16)
    draft = (target.float() + 0.7 * torch.randn(B, L, V, device=dev, generator=g)).to(torch.bfloat16)
    little = (draft.float() + 0.7 * torch.randn(B, L, V, device=dev, generator=g)).to(torch.bfloat16)
    x = torch.randint(0, V, (B, L), device=dev, generator=g)
    return little, draft, target, x


def run_c4(args):
    """BASELINE config 4: tridecoding's two verification stages through the mirrored operators - stage 1: the
    little model's gamma2=4 proposals against the draft model, stage 2: up to 9 positions against the target,
    both with transfer_top_k=1024 (rebuild_topk_uniform_probs -> build_draft_probs_override ->
    resolve_stage_verification, /root/reference/src/baselines.py:1426-1678, src/decoding_ops.py:353-397)."""
    from duodecoding_b200 import decoding_ops, model_gpu, proposal_utils, utils
    import duodecoding_b200.ops as ops

    world, rank, local = dist_env()
    if rank != 0:
        return 0
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the verification path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    B, V, P, g2, g1, k, T = (C4[n] for n in ("B", "V", "P", "gamma2", "gamma1", "k", "T"))
    little, draft, target, x = _c4_logits(dev, 77)
    caches = []
    for hist, rows in ((little, P - 1 + g2), (draft, P - 1 + g1), (target, P + g1)):
        c = model_gpu.KVCacheModel(_NoModel(V, dev), T, 0, 0.0)
        c._store_rows(hist[:, :rows], 0)
        c._past_key_values = _Len(rows)
        caches.append((c, rows))

    def step():
        for c, rows in caches:  # new rows arrive every step: their probabilities are materialised again
            c._current_seq_len, c._past_key_values.n, c._prob_done = rows, rows, P - 1
        (lc, _), (dc, _), (tc, _) = caches
        rb = utils.rebuild_topk_uniform_probs(lc.prob_history[:, P - 1:P - 1 + g2, :], k)
        ov = proposal_utils.build_draft_probs_override(lc, P, rb)
        dc._current_seq_len = dc._past_key_values.n = P + g2  # stage 1 sees the draft model one step ahead
        a1 = decoding_ops.resolve_stage_verification(lc, dc, x, P, g2, output_device=dev, draft_probs_override=ov)
        dc._current_seq_len = dc._past_key_values.n = P - 1 + g1
        rb2 = utils.rebuild_topk_uniform_probs(dc.prob_history[:, P - 1:P - 1 + g1, :], k)
        ov2 = proposal_utils.build_draft_probs_override(dc, P, rb2)
        a2 = decoding_ops.resolve_stage_verification(dc, tc, x, P, g1, output_device=dev, draft_probs_override=ov2)
        return a1[0], a2[0]

    K, W = max(10, min(args.steps, 200)), max(args.warmup, 3)
    for _ in range(W):
        step()
    torch.cuda.synchronize()
    clocks = ClockSampler(local)
    t0 = time.perf_counter()
    for _ in range(K):
        step()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / K
    clk = clocks.stop()
    # the dominant kernel alone: transfer_top_k rebuild of the 9 stage-2 rows
    dcm = caches[1][0]  # (the step's rollback shortened the history: restore the 9 stage-2 rows first)
    dcm._current_seq_len = dcm._past_key_values.n = P - 1 + g1
    dcm._prob_done = P - 1
    probs = dcm.prob_history[:, P - 1:P - 1 + g1, :].contiguous()
    assert probs.shape == (B, g1, V), probs.shape
    for _ in range(3):
        ops.topk_rebuild(probs, k, 0)
    torch.cuda.synchronize()
    side = torch.cuda.Stream(device=dev)  # (the host side of the op costs more than the kernel: time a graph of launches)
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr, stream=side):
        for _ in range(50):
            ops.topk_rebuild(probs, k, 0)
    gr.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    gr.replay()
    e1.record()
    torch.cuda.synchronize()
    kus = e0.elapsed_time(e1) * 1e3 / 50
    kbytes = 2 * probs.numel() * 2
    peak, peak_src = hbm_peak()
    positions = B * ((g2 + 1) + (g1 + 1))
    rows_norm = g2 + (g1 + 1) + (g1 + 1)  # probability rows materialised per step (little, draft, target)
    alg = B * V * 2 * (2 * rows_norm + 2 * (g2 + g1) + 4)
    cpu = None
    if not args.no_cpu:
        cpu = c4_cpu(little, draft, target, x)
    line = {
        "metric": METRIC, "value": positions / dt, "unit": UNIT, "n_gpus": 1, "steps": K, "warmup": W,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic",
        "config": {"workload": "C4 two-stage tridecoding verification: gamma2=4 then 9 positions, V=32000, bf16, "
                               "transfer_top_k=1024, through the mirrored operators", **C4,
                   "api": "utils.rebuild_topk_uniform_probs + proposal_utils.build_draft_probs_override + "
                          "decoding_ops.resolve_stage_verification (probability-space operators)",
                   "timing": "host clock around K steps + synchronize (the operators read Python ints back)",
                   "l2": "working set 3 models x 18 rows x B x V x 2 B = 27 MB: L2 resident, as in the decode loop"},
        "clocks": clk, "gpu_launches": 12 * K,
        "e2e": {"value": positions / dt, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 16,
                "note": "histories live on the device (the models write them); the read-back is the two accepted counts"},
        "roofline": {"bound": "hbm", "achieved": kbytes / (kus * 1e-6) / 1e9, "peak": peak, "unit": "GB/s",
                     "frac": kbytes / (kus * 1e-6) / 1e9 / peak, "traffic": None,
                     "kernel": "rowsel::topk_reg_kernel<__nv_bfloat16> mode 0, [8, 9, 32000], k=1024 (graph of 50 launches)",
                     "alg_bytes_per_launch": kbytes, "us_per_launch": kus, "peak_source": peak_src,
                     "step_alg_bytes": alg, "step_GBps": alg / dt / 1e9},
        "cpu_baseline": cpu,
    }
    emit(line)
    return 0


def c4_cpu(little, draft, target, x):
    """The same two stages on the oracle port (torch CPU, all host threads)."""
    import oracle

    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    B, V, P, g2, g1, k, T = (C4[n] for n in ("B", "V", "P", "gamma2", "gamma1", "k", "T"))
    l, d, t, xs = little.cpu().float(), draft.cpu().float(), target.cpu().float(), x.cpu()

    def stage(ql, pl, gam):
        q = oracle.rebuild_topk_uniform_probs(oracle.norm_logits(ql[:, P - 1:P - 1 + gam], T, 0, 0.0), k)
        p = oracle.norm_logits(pl[:, P - 1:P + gam], T, 0, 0.0)
        r = torch.rand(B, gam)
        acc = oracle.compute_acceptance(q, p[:, :gam], xs[:, P:P + gam], P, r)
        a = acc["accepted_count"]
        noise = torch.empty(B, V).exponential_(1.0)
        return oracle.sample_accept_token(p[:, gam], noise) if a == gam else oracle.sample_reject_token(p[:, a], q[:, a], noise)

    def step():
        stage(l, d, g2)
        stage(d, t, g1)

    step()
    n = 10
    t0 = time.perf_counter()
    for _ in range(n):
        step()
    dt = (time.perf_counter() - t0) / n
    return {"value": B * ((g2 + 1) + (g1 + 1)) / dt, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"{n} steps of the full C4 step ({dt * 1e3:.1f} ms/step)"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=["c2", "c3", "c4", "c5"],
                    help="c2: BASELINE config 2 (default, batch per rank); c3: B=32 bf16 (--temperature 0|1); "
                         "c4: two-stage tridecoding verification; c5: vocabulary-sharded config 5")
    ap.add_argument("--temperature", type=float, default=None, help="c3 only: 0 (greedy) or 1")
    ap.add_argument("--top-p", type=float, default=0.0, help="c5 only: nucleus filter on both models (sharded top-p)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-check", action="store_true", help="skip the oracle check of one step before timing")
    ap.add_argument("--no-c5", action="store_true", help="N > 1: skip the vocabulary-sharded c5 sub-record")
    ap.add_argument("--clock-seconds", type=float, default=0.6,
                    help="untimed replays of the same graph before and after the timed one, for the clock sampler")
    args = ap.parse_args()
    # stdout carries ONE JSON line: everything a library prints there meanwhile (the NCCL version banner of either
    # rank, for one) goes to stderr; the JSON line is written to the saved descriptor at the end
    global _STDOUT
    sys.stdout.flush()
    _STDOUT = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        return run_reference(args)
    if args.workload == "c5":
        return run_c5(args)
    if args.workload == "c4":
        return run_c4(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
