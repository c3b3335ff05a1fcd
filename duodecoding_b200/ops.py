"""Tensor-level entry points: torch CUDA tensors in, C-ABI calls out.

torch is plumbing here (device memory, the current stream); all arithmetic
happens in the sm_100a kernels of ``libduodec.so``.  Every function raises on
CPU tensors - there is no fallback path.
"""

from __future__ import annotations

from collections import OrderedDict
from typing import Dict, Optional, Tuple

import torch

from . import _lib
from ._lib import BF16, F32, check, lib

_DTYPES = {torch.float32: F32, torch.bfloat16: BF16}
# verify workspaces: one per (device, stream, B, gamma) - a workspace holds the records of ONE launch in
# flight, so launches on different streams must not share it; least recently used entries are dropped
_WORKSPACES: "OrderedDict[Tuple[int, int, int, int], torch.Tensor]" = OrderedDict()
_WORKSPACE_LIMIT_BYTES = 1 << 30


def _need_cuda(*tensors: torch.Tensor) -> None:
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError(
                "duodecoding_b200 ops run on CUDA tensors only (no CPU fallback); got a tensor on " + str(t.device)
            )


def _code(t: torch.Tensor) -> int:
    try:
        return _DTYPES[t.dtype]
    except KeyError:
        raise TypeError(f"unsupported dtype {t.dtype}; float32 and bfloat16 are implemented") from None


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _rows(t: torch.Tensor) -> torch.Tensor:
    """View [..., V] as [rows, V] with unit vocab stride (copy only when needed)."""
    V = t.shape[-1]
    t2 = t.reshape(-1, V)
    if t2.numel() and t2.stride(-1) != 1:
        t2 = t2.contiguous()
    return t2


def _lastdim_contig(t: torch.Tensor) -> torch.Tensor:
    return t if (t.numel() == 0 or t.stride(-1) == 1) else t.contiguous()


# --------------------------------------------------------------------------
# torch.library registration: every kernel launch of this package is a `duodec::` custom op with its outputs
# annotated as mutated (`Tensor(a!)`), implemented for CUDA by the C-ABI call (ctypes) and for Meta by a
# no-op - the "thin torch custom-op layer" over libduodec.so.  The functions below prepare views / outputs
# and then call torch.ops.duodec.*; nothing else in the package launches kernels.
# --------------------------------------------------------------------------
_TL = torch.library.Library("duodec", "DEF")


def _register(schema: str, fn) -> None:
    name = schema.split("(")[0]
    _TL.define(schema)
    _TL.impl(name, fn, "CUDA")
    _TL.impl(name, lambda *a, **k: None, "Meta")


def _op_norm_logits(logits, out, temperature, top_k, top_p):
    V = logits.shape[-1]
    with torch.cuda.device(logits.device):
        if logits.dim() == 3:
            check(lib.duodec_norm_logits_batched(
                logits.data_ptr(), _code(logits), logits.stride(0), logits.stride(1), out.data_ptr(), _code(out),
                out.stride(0), out.stride(1), logits.shape[0], logits.shape[1], V, float(temperature), int(top_k),
                float(top_p), _stream()), "duodec_norm_logits_batched")
        else:
            check(lib.duodec_norm_logits(
                logits.data_ptr(), _code(logits), logits.stride(0), out.data_ptr(), _code(out), out.stride(0),
                logits.shape[0], V, float(temperature), int(top_k), float(top_p), _stream()), "duodec_norm_logits")


def _op_verify_fused(dl, tl, tokens, r, noise, T, k, p, dT, dk, dp, flags, accept_mask, sel_p, sel_q, accepted_len,
                     token, all_accepted, ws):
    B, gamma = tokens.shape
    with torch.cuda.device(tl.device):
        check(lib.duodec_verify_fused(
            dl.data_ptr(), tl.data_ptr(), _code(tl), dl.stride(0), dl.stride(1), tl.stride(0), tl.stride(1),
            B, gamma, dl.shape[-1], tl.shape[-1],
            tokens.data_ptr(), tokens.stride(0) if gamma else 0, r.data_ptr(), r.stride(0) if gamma else 0,
            noise.data_ptr(), noise.stride(0), float(T), int(k), float(p), float(dT), int(dk), float(dp), int(flags),
            accept_mask.data_ptr(), sel_p.data_ptr(), sel_q.data_ptr(), accepted_len.data_ptr(), token.data_ptr(),
            all_accepted.data_ptr(), ws.data_ptr(), ws.numel(), _stream()), "duodec_verify_fused")


def _op_accept_probs(dq, tp, tokens, r, V, accept_mask, sel_q, sel_p, accepted_len):
    B, gamma = tokens.shape
    with torch.cuda.device(dq.device):
        check(lib.duodec_accept_probs(
            dq.data_ptr(), dq.stride(0), dq.stride(1) if gamma else 0, tp.data_ptr(), tp.stride(0),
            tp.stride(1) if gamma else 0, _code(dq), tokens.data_ptr(), tokens.stride(0) if gamma else 0,
            r.data_ptr(), r.stride(0) if gamma else 0, B, gamma, int(V), accept_mask.data_ptr(), sel_q.data_ptr(),
            sel_p.data_ptr(), accepted_len.data_ptr(), _stream()), "duodec_accept_probs")


def _op_sample_rows(p, nz, out):
    with torch.cuda.device(p.device):
        check(lib.duodec_sample_rows(p.data_ptr(), _code(p), p.stride(0), nz.data_ptr(), nz.stride(0), p.shape[0],
                                     p.shape[1], out.data_ptr(), _stream()), "duodec_sample_rows")


def _op_residual_sample(tp, dq, nz, out):
    with torch.cuda.device(tp.device):
        check(lib.duodec_residual_sample(tp.data_ptr(), tp.stride(0), dq.data_ptr(), dq.stride(0), _code(tp),
                                         nz.data_ptr(), nz.stride(0), tp.shape[0], tp.shape[1], out.data_ptr(),
                                         _stream()), "duodec_residual_sample")


def _op_max_fn(x2, out):
    with torch.cuda.device(x2.device):
        check(lib.duodec_max_fn(x2.data_ptr(), _code(x2), x2.stride(0), x2.shape[0], x2.shape[1], out.data_ptr(),
                                out.stride(0), _stream()), "duodec_max_fn")


def _op_topk_rebuild(p2, k, mode, out):
    with torch.cuda.device(p2.device):
        check(lib.duodec_topk_rebuild(p2.data_ptr(), _code(p2), p2.stride(0), out.data_ptr(), out.stride(0),
                                      p2.shape[0], p2.shape[1], int(k), int(mode), _stream()), "duodec_topk_rebuild")


def _op_kv_truncate(seq_len, end_vec, end_pos):
    with torch.cuda.device(seq_len.device):
        check(lib.duodec_kv_truncate(seq_len.data_ptr(), end_vec.data_ptr() if end_vec is not None else None,
                                     int(end_pos), seq_len.numel(), _stream()), "duodec_kv_truncate")


_register("norm_logits(Tensor logits, Tensor(a!) out, float temperature, int top_k, float top_p) -> ()", _op_norm_logits)
_register("verify_fused(Tensor draft, Tensor target, Tensor tokens, Tensor r, Tensor noise, float temperature, "
          "int top_k, float top_p, float draft_temperature, int draft_top_k, float draft_top_p, int flags, "
          "Tensor(a!) accept_mask, Tensor(b!) sel_target_p, Tensor(c!) sel_draft_p, Tensor(d!) accepted_len, "
          "Tensor(e!) token, Tensor(f!) all_accepted, Tensor(g!) workspace) -> ()", _op_verify_fused)
_register("accept_probs(Tensor draft_probs, Tensor target_probs, Tensor tokens, Tensor r, int V, "
          "Tensor(a!) accept_mask, Tensor(b!) sel_draft_p, Tensor(c!) sel_target_p, Tensor(d!) accepted_len) -> ()",
          _op_accept_probs)
_register("sample_rows(Tensor probs, Tensor noise, Tensor(a!) out) -> ()", _op_sample_rows)
_register("residual_sample(Tensor target_probs, Tensor draft_probs, Tensor noise, Tensor(a!) out) -> ()",
          _op_residual_sample)
_register("max_fn(Tensor x, Tensor(a!) out) -> ()", _op_max_fn)
_register("topk_rebuild(Tensor probs, int k, int mode, Tensor(a!) out) -> ()", _op_topk_rebuild)
_register("kv_truncate(Tensor(a!) seq_len, Tensor? end_pos_vec, int end_pos) -> ()", _op_kv_truncate)
_D = torch.ops.duodec


def set_team_size(team: int) -> None:
    """Tuning hook: force the thread-block-cluster size (0 = automatic)."""
    lib.duodec_set_team_size(int(team))


FAST_PATH = 5  # row-pair scan kernel + sliced sample kernel
_PATH = 0       # DUODEC_VERIFY_PATH: 0 automatic, 1 general team kernel, 5 row-pair kernel
_PHASES = 0     # DUODEC_VERIFY_PHASES: 0 / 3 both kernels, 1 scan only, 2 sample only
ROW0_RULE = 1   # DUODEC_VERIFY_ROW0_RULE


def set_stream_path(path) -> None:
    """Test hook: device path of the following verify_fused calls.  True: automatic (the row-capture fast
    path when the call qualifies); False / 0: the general team kernel; 5: row-pair scan kernel.  The choice travels in the call's ``flags`` argument - the library keeps no path state."""
    global _PATH
    if isinstance(path, bool):
        _PATH = 0 if path else 1
    else:
        _PATH = 1 if int(path) == 0 else int(path)


def set_phase_mask(mask: int) -> None:
    """Measurement hook: 1 = scan kernel only, 2 = sample kernel only, 3 = both (default)."""
    global _PHASES
    _PHASES = int(mask) & 3


def set_stats_buffer(buf: Optional[torch.Tensor]) -> None:
    """Tuning hook: int32[32] CUDA tensor the row-stream kernel counts into (None disables)."""
    if buf is not None and (not buf.is_cuda or buf.dtype != torch.int32 or buf.numel() < 32):
        raise TypeError("stats buffer must be a CUDA int32 tensor with >= 32 entries")
    lib.duodec_set_stats_buffer(buf.data_ptr() if buf is not None else None)


def team_size(V: int, rows_resident: int = 1) -> int:
    return int(lib.duodec_team_size(int(V), int(rows_resident)))


# --------------------------------------------------------------------------
def norm_logits(
    logits: torch.Tensor,
    temperature: float,
    top_k: int = 0,
    top_p: float = 0.0,
    out: Optional[torch.Tensor] = None,
    out_dtype: Optional[torch.dtype] = None,
) -> torch.Tensor:
    """softmax(top_k_top_p_filter(logits / T)), or the argmax one-hot at T == 0."""
    _need_cuda(logits, out)
    V = logits.shape[-1]
    if out is None:
        if out_dtype is None:
            out_dtype = torch.float32 if temperature == 0 else logits.dtype
        out = torch.empty(logits.shape, dtype=out_dtype, device=logits.device)
    if logits.dim() == 3 and out.dim() == 3 and logits.stride(-1) == 1 and out.stride(-1) == 1 and \
            logits.shape == out.shape:
        # [B, S, V] views with their own batch / row strides (history buffers): one launch, no copy
        if logits.shape[0] and logits.shape[1]:
            _D.norm_logits(logits, out, float(temperature), int(top_k), float(top_p))
        return out
    x = _rows(logits)
    rows = x.shape[0]
    o = out.reshape(-1, V) if out.dim() != 2 else out
    if o.numel() and (o.stride(-1) != 1 or o.data_ptr() != out.data_ptr()):
        raise ValueError("out must have unit stride along the vocabulary and rows that flatten without a copy")
    if rows:
        _D.norm_logits(x, o, float(temperature), int(top_k), float(top_p))
    return out


def _workspace(device: torch.device, B: int, gamma: int) -> torch.Tensor:
    idx = device.index if device.index is not None else torch.cuda.current_device()
    stream = torch.cuda.current_stream(device)
    key = (idx, stream.cuda_stream, B, gamma)
    ws = _WORKSPACES.get(key)
    if ws is not None:
        _WORKSPACES.move_to_end(key)
        return ws
    nbytes = int(lib.duodec_verify_workspace_bytes(B, gamma))
    with torch.cuda.stream(stream):
        ws = torch.zeros(nbytes, dtype=torch.uint8, device=device)  # zero-filled on the stream that will use it
    _WORKSPACES[key] = ws
    total = sum(int(t.numel()) for t in _WORKSPACES.values())
    while total > _WORKSPACE_LIMIT_BYTES and len(_WORKSPACES) > 1:
        _, old = _WORKSPACES.popitem(last=False)
        total -= int(old.numel())
    return ws


def verify_fused(
    draft_logits: torch.Tensor,  # [B, >=gamma, Vd]
    target_logits: torch.Tensor,  # [B, >=gamma+1, Vt]
    tokens: torch.Tensor,  # [B, gamma] int64
    r: torch.Tensor,  # [B, gamma] float32
    noise: torch.Tensor,  # [B, >=max(Vd,Vt)] float32 Exp(1)
    temperature: float = 1.0,
    top_k: int = 0,
    top_p: float = 0.0,
    draft_temperature: Optional[float] = None,
    draft_top_k: Optional[int] = None,
    draft_top_p: Optional[float] = None,
    out: Optional[Dict[str, torch.Tensor]] = None,
    row0_rule: bool = False,
) -> Dict[str, torch.Tensor]:
    """One fused verification step on logits for every sequence of the batch.  ``row0_rule``: the
    reference's batch rule - sequence 0's accepted count positions every sequence (fast paths only)."""
    _need_cuda(draft_logits, target_logits, tokens, r, noise)
    B, gamma = tokens.shape
    dev = target_logits.device
    if target_logits.dtype != draft_logits.dtype:
        raise TypeError("draft and target logits must share a dtype")
    if target_logits.shape[1] < gamma + 1 or (gamma > 0 and draft_logits.shape[1] < gamma):
        raise ValueError("need gamma draft rows and gamma+1 target rows")
    dl = _lastdim_contig(draft_logits)
    tl = _lastdim_contig(target_logits)
    tokens = _lastdim_contig(tokens.to(torch.int64))
    r = _lastdim_contig(r.to(torch.float32))
    noise = _lastdim_contig(noise.to(torch.float32))
    Vd, Vt = dl.shape[-1], tl.shape[-1]
    if noise.shape[-1] < max(Vd, Vt) or noise.shape[0] != B:
        raise ValueError("noise must be [B, >= max(Vd, Vt)]")
    dT = temperature if draft_temperature is None else draft_temperature
    dk = top_k if draft_top_k is None else draft_top_k
    dp = top_p if draft_top_p is None else draft_top_p
    if out is None:
        out = {
            "accept_mask": torch.empty((B, gamma), dtype=torch.bool, device=dev),
            "selected_target_p": torch.empty((B, gamma), dtype=torch.float32, device=dev),
            "selected_draft_p": torch.empty((B, gamma), dtype=torch.float32, device=dev),
            "accepted_len": torch.empty((B,), dtype=torch.int64, device=dev),
            "token": torch.empty((B,), dtype=torch.int64, device=dev),
            "all_accepted": torch.empty((B,), dtype=torch.bool, device=dev),
        }
    if B == 0:
        return out
    ws = _workspace(dev, B, gamma)
    _D.verify_fused(dl, tl, tokens, r, noise, float(temperature), int(top_k), float(top_p), float(dT), int(dk), float(dp),
                    (ROW0_RULE if row0_rule else 0) | ((_PATH & 15) << 4) | ((_PHASES & 3) << 8),
                    out["accept_mask"], out["selected_target_p"], out["selected_draft_p"], out["accepted_len"],
                    out["token"], out["all_accepted"], ws)
    return out


def accept_probs(
    draft_probs: torch.Tensor,  # [B, gamma, V]
    target_probs: torch.Tensor,  # [B, gamma, V]
    tokens: torch.Tensor,  # [B, gamma]
    r: torch.Tensor,  # [B, gamma]
) -> Dict[str, torch.Tensor]:
    """Gather p, q at the drafted tokens, ``r <= p/q``, leading accepted run per sequence."""
    _need_cuda(draft_probs, target_probs, tokens, r)
    B, gamma = tokens.shape
    dev = draft_probs.device
    if target_probs.device != dev:
        target_probs = target_probs.to(dev)
    if draft_probs.dtype != target_probs.dtype or draft_probs.dtype not in _DTYPES:
        draft_probs, target_probs = draft_probs.float(), target_probs.float()
    dq = _lastdim_contig(draft_probs)
    tp = _lastdim_contig(target_probs)
    tokens = _lastdim_contig(tokens.to(device=dev, dtype=torch.int64))
    r = _lastdim_contig(r.to(device=dev, dtype=torch.float32))
    V = min(dq.shape[-1], tp.shape[-1])
    out = {
        "accept_mask": torch.empty((B, gamma), dtype=torch.bool, device=dev),
        "selected_draft_p": torch.empty((B, gamma), dtype=dq.dtype, device=dev),
        "selected_target_p": torch.empty((B, gamma), dtype=dq.dtype, device=dev),
        "accepted_len": torch.empty((B,), dtype=torch.int64, device=dev),
    }
    if B == 0:
        return out
    _D.accept_probs(dq, tp, tokens, r, V, out["accept_mask"], out["selected_draft_p"], out["selected_target_p"],
                    out["accepted_len"])
    return out


def draw_noise(rows: int, V: int, device: torch.device, generator: Optional[torch.Generator] = None) -> torch.Tensor:
    """Exp(1) draws for the race sampler - the same draw torch.multinomial makes internally."""
    return torch.empty((rows, V), dtype=torch.float32, device=device).exponential_(1.0, generator=generator)


def sample_rows(probs: torch.Tensor, noise: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``sample(probs)`` -> [rows, 1] int64."""
    _need_cuda(probs, noise)
    p = _rows(probs)
    if p.dtype not in _DTYPES:
        p = p.float()
    rows, V = p.shape
    if noise is None:
        noise = draw_noise(rows, V, p.device)
    nz = _rows(noise.to(torch.float32))
    out = torch.empty((rows,), dtype=torch.int64, device=p.device)
    if rows:
        _D.sample_rows(p, nz, out)
    return out.reshape(*probs.shape[:-1], 1)


def residual_sample(target_probs: torch.Tensor, draft_probs: torch.Tensor,
                    noise: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``sample(max_fn(target_probs - draft_probs))`` -> [rows, 1] int64."""
    _need_cuda(target_probs, draft_probs, noise)
    if target_probs.dtype != draft_probs.dtype or target_probs.dtype not in _DTYPES:
        target_probs, draft_probs = target_probs.float(), draft_probs.float()
    tp = _rows(target_probs)
    dq = _rows(draft_probs)
    rows, V = tp.shape
    if dq.shape != tp.shape:
        raise ValueError("target and draft rows must have the same shape")
    if noise is None:
        noise = draw_noise(rows, V, tp.device)
    nz = _rows(noise.to(torch.float32))
    out = torch.empty((rows,), dtype=torch.int64, device=tp.device)
    if rows:
        _D.residual_sample(tp, dq, nz, out)
    return out.reshape(*target_probs.shape[:-1], 1)


def max_fn(x: torch.Tensor) -> torch.Tensor:
    """``norm(max(x, 0))`` over the last dim of a 2-D tensor, fp32 out."""
    _need_cuda(x)
    if x.dtype not in _DTYPES:
        x = x.float()
    x2 = _rows(x)
    rows, V = x2.shape
    out = torch.empty((rows, V), dtype=torch.float32, device=x.device)
    if rows:
        _D.max_fn(x2, out)
    return out.reshape(x.shape)


def topk_rebuild(probs: torch.Tensor, k: int, mode: int) -> torch.Tensor:
    """transfer_top_k compression family (modes in include/duodec.h)."""
    _need_cuda(probs)
    p = probs if probs.dtype in _DTYPES else probs.float()
    p2 = _rows(p)
    rows, V = p2.shape
    out = torch.empty((rows, V), dtype=p2.dtype, device=p2.device)
    if rows:
        _D.topk_rebuild(p2, int(k), int(mode), out)
    return out.reshape(probs.shape).to(probs.dtype)


def kv_truncate(seq_len: torch.Tensor, end_pos) -> torch.Tensor:
    """In-place ``seq_len[b] = min(seq_len[b], end_pos[b])`` on an int32 length vector."""
    _need_cuda(seq_len)
    if seq_len.dtype != torch.int32 or not seq_len.is_contiguous():
        raise TypeError("seq_len must be a contiguous int32 tensor")
    B = seq_len.numel()
    vec = None
    scalar = 0
    if isinstance(end_pos, torch.Tensor):
        vec = end_pos.to(device=seq_len.device, dtype=torch.int64).contiguous()
        if vec.numel() != B:
            raise ValueError("end_pos tensor must have one entry per sequence")
    else:
        scalar = int(end_pos)
    if B:
        _D.kv_truncate(seq_len, vec, scalar)
    return seq_len
