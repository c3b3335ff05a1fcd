"""ctypes binding of ``libduodec.so`` (the C ABI declared in include/duodec.h).

There is no fallback: if the shared library is missing or a symbol is absent
the import fails loudly, and every op raises on non-CUDA tensors.
"""

from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DUODEC_LIB") or os.path.join(_HERE, "libduodec.so")  # env: tuning builds only

F32 = 0
BF16 = 1

c_void_p = ctypes.c_void_p
c_int = ctypes.c_int
c_i64 = ctypes.c_int64
c_float = ctypes.c_float
c_size_t = ctypes.c_size_t

# name -> (restype, argtypes); mirrors include/duodec.h one to one
SIGNATURES = {
    "duodec_version": (ctypes.c_char_p, []),
    "duodec_team_size": (c_int, [c_int, c_int]),
    "duodec_set_team_size": (None, [c_int]),
    "duodec_set_stats_buffer": (None, [c_void_p]),
    "duodec_norm_logits": (
        c_int,
        [c_void_p, c_int, c_i64, c_void_p, c_int, c_i64, c_int, c_int, c_float, c_int, c_float, c_void_p],
    ),
    "duodec_norm_logits_batched": (
        c_int,
        [c_void_p, c_int, c_i64, c_i64, c_void_p, c_int, c_i64, c_i64, c_int, c_int, c_int, c_float, c_int, c_float,
         c_void_p],
    ),
    "duodec_verify_workspace_bytes": (c_size_t, [c_int, c_int]),
    "duodec_verify_fused": (
        c_int,
        [c_void_p, c_void_p, c_int, c_i64, c_i64, c_i64, c_i64, c_int, c_int, c_int, c_int]  # rows / shapes
        + [c_void_p, c_i64, c_void_p, c_i64, c_void_p, c_i64]  # tokens, r, noise
        + [c_float, c_int, c_float, c_float, c_int, c_float]  # target warp, draft warp
        + [c_int]  # flags
        + [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]  # outputs
        + [c_void_p, c_size_t, c_void_p],  # workspace, stream
    ),
    "duodec_token_ranks": (c_int, [c_void_p, c_int, c_i64, c_int, c_int, c_void_p, c_i64, c_void_p, c_void_p]),
    "duodec_verify_static_max_bytes": (c_size_t, []),
    "duodec_verify_fused_static": (
        c_int,
        [c_void_p, c_void_p, c_int, c_i64, c_i64, c_i64, c_i64, c_int, c_int, c_int, c_int]  # rows / shapes
        + [c_void_p, c_i64, c_void_p, c_i64, c_void_p, c_i64]  # tokens, r, noise
        + [c_float, c_int, c_float, c_float, c_int, c_float]  # target warp, draft warp
        + [c_int]  # flags
        + [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]  # outputs
        + [c_void_p],  # stream
    ),
    "duodec_accept_probs": (
        c_int,
        [c_void_p, c_i64, c_i64, c_void_p, c_i64, c_i64, c_int, c_void_p, c_i64, c_void_p, c_i64]
        + [c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    ),
    "duodec_sample_rows": (c_int, [c_void_p, c_int, c_i64, c_void_p, c_i64, c_int, c_int, c_void_p, c_void_p]),
    "duodec_residual_sample": (
        c_int,
        [c_void_p, c_i64, c_void_p, c_i64, c_int, c_void_p, c_i64, c_int, c_int, c_void_p, c_void_p],
    ),
    "duodec_max_fn": (c_int, [c_void_p, c_int, c_i64, c_int, c_int, c_void_p, c_i64, c_void_p]),
    "duodec_topk_rebuild": (c_int, [c_void_p, c_int, c_i64, c_void_p, c_i64, c_int, c_int, c_int, c_int, c_void_p]),
    "duodec_kv_truncate": (c_int, [c_void_p, c_void_p, c_i64, c_int, c_void_p]),
    "duodec_uncertainty_tokens": (
        c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p, c_i64, c_int, c_void_p, c_void_p]),
    "duodec_compressed_vocab_size": (
        c_int, [c_void_p, c_void_p, c_int, ctypes.c_double, ctypes.c_double, c_void_p, c_void_p]),
    "duodec_compact_nonzero": (
        c_int, [c_void_p, c_int, c_i64, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "duodec_shard_topk_apply": (
        c_int,
        [c_void_p, c_int, c_i64, c_void_p, c_i64, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
         c_void_p]),
    "duodec_shard_stats": (
        c_int,
        [c_void_p, c_void_p, c_int, c_i64, c_i64, c_i64, c_i64, c_int, c_int, c_int, c_int, c_void_p, c_i64]
        + [c_float, c_float, c_void_p, c_void_p, c_void_p],
    ),
    "duodec_shard_accept": (
        c_int,
        [c_void_p, c_int, c_int, c_int, c_void_p, c_i64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
         c_void_p, c_void_p, c_i64, c_void_p, c_void_p],
    ),
    "duodec_shard_sample": (
        c_int,
        [c_void_p, c_void_p, c_int, c_i64, c_i64, c_i64, c_i64, c_int, c_int, c_int, c_int, c_float, c_float]
        + [c_void_p, c_void_p, c_i64, c_int, c_void_p, c_void_p, c_int, c_void_p],
    ),
    "duodec_shard_nucleus_levels": (c_int, [c_int, c_float, c_float]),
    "duodec_shard_nucleus_level": (
        c_int,
        [c_void_p, c_void_p, c_int, c_i64, c_i64, c_i64, c_i64, c_int, c_int, c_int, c_int, c_float, c_float]
        + [c_float, c_float, c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p],
    ),
    "duodec_shard_nucleus_reduce": (
        c_int, [c_void_p, c_int, c_int, c_int, c_int, c_float, c_float, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "duodec_shard_nucleus_tie": (
        c_int,
        [c_void_p, c_void_p, c_int, c_i64, c_i64, c_i64, c_i64, c_int, c_int, c_int, c_int, c_float, c_float]
        + [c_void_p, c_void_p, c_int, c_void_p, c_void_p],
    ),
    "duodec_shard_nucleus_close": (
        c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "duodec_shard_finish": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
}


class ShardXchg(ctypes.Structure):
    """duodec_shard_xchg (include/duodec.h): the fused peer-memory exchange of the sharded step."""

    _fields_ = [("peers", c_void_p), ("rank", c_int), ("world", c_int), ("off_stats", c_i64), ("off_partial", c_i64),
                ("mode", c_int)]


class DuodecError(RuntimeError):
    """A C-ABI call returned non-zero (negative: argument error, positive: cudaError_t)."""


def _load() -> ctypes.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc -gencode arch=compute_100a,code=sm_100a).  duodecoding_b200 has no CPU or eager fallback."
        )
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing: intended
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()

_ERRORS = {-1: "DUODEC_EINVAL (bad shape, null pointer or dtype)", -2: "DUODEC_ETOOBIG (row too large)",
           -3: "DUODEC_EWORKSPACE (workspace too small)",
           -4: "DUODEC_ESTATICMAX (the call needs more scratch than the library's built-in 12 MiB array: use "
               "duodec_verify_fused with a caller workspace)"}


def check(code: int, what: str) -> None:
    if code == 0:
        return
    if code < 0:
        raise DuodecError(f"{what}: {_ERRORS.get(code, code)}")
    raise DuodecError(f"{what}: cudaError_t {code}")


def version() -> str:
    return lib.duodec_version().decode()
