"""duodec_verify_fused_static: the no-allocation variant of the fused verification (scratch in a static device
array inside the library, include/duodec.h) against duodec_verify_fused, through the C ABI.  Collected last."""
import ctypes

import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu

NUCLEUS_RTOL = 5e-4  # a top-p boundary entry may flip with the exponent-offset hint (tests/test_gpu_parity.py)


def _case(seed, B, gamma, V, dtype=torch.float32, top_p=0.9):
    g = torch.Generator().manual_seed(seed)
    tl = (torch.randn(B, gamma + 1, V, generator=g) * 3).to(dtype)
    dl = (tl[:, :gamma].float() + torch.randn(B, gamma, V, generator=g)).to(dtype)
    q = oracle.norm_logits(dl.float(), 1.0, 0, top_p)
    toks = torch.stack(
        [oracle.sample(q[:, j], torch.empty(B, V).exponential_(1, generator=g))[:, 0] for j in range(gamma)], 1)
    r = torch.rand(B, gamma, generator=g)
    noise = torch.empty(B, V).exponential_(1, generator=g)
    return dl, tl, toks, r, noise


def _outputs(B, gamma, dev):
    return {
        "accept_mask": torch.zeros((B, gamma), dtype=torch.bool, device=dev),
        "selected_target_p": torch.zeros((B, gamma), dtype=torch.float32, device=dev),
        "selected_draft_p": torch.zeros((B, gamma), dtype=torch.float32, device=dev),
        "accepted_len": torch.full((B,), -1, dtype=torch.int64, device=dev),
        "token": torch.full((B,), -1, dtype=torch.int64, device=dev),
        "all_accepted": torch.zeros((B,), dtype=torch.bool, device=dev),
    }


def _call_static(lib, dl, tl, toks, r, noise, T, top_k, top_p, out):
    """The raw C-ABI call: pointers, strides in elements, the current stream; nothing is allocated in here."""
    B, gamma = toks.shape
    dtype = 0 if tl.dtype == torch.float32 else 1
    stream = torch.cuda.current_stream(tl.device).cuda_stream
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    return lib.duodec_verify_fused_static(
        p(dl), p(tl), dtype, dl.stride(0), dl.stride(1), tl.stride(0), tl.stride(1), B, gamma, dl.shape[-1], tl.shape[-1],
        p(toks), toks.stride(0), p(r), r.stride(0), p(noise), noise.stride(0),
        T, top_k, top_p, T, top_k, top_p, 0,
        p(out["accept_mask"]), p(out["selected_target_p"]), p(out["selected_draft_p"]), p(out["accepted_len"]),
        p(out["token"]), p(out["all_accepted"]), ctypes.c_void_p(stream))


def _same(got, want, what):
    for k in ("accepted_len", "token", "accept_mask", "all_accepted"):
        assert torch.equal(got[k], want[k]), (what, k)
    for k in ("selected_target_p", "selected_draft_p"):
        torch.testing.assert_close(got[k], want[k], rtol=NUCLEUS_RTOL, atol=1e-9, msg=lambda m: f"{what} {k}: {m}")


def test_static_workspace_entry_point_matches_verify_fused_on_consecutive_calls():
    import duodecoding_b200.ops as ops
    from duodecoding_b200 import _lib

    lib = _lib.lib
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    assert lib.duodec_verify_static_max_bytes() == 12 << 20
    # two calls in a row with the same shape (the second runs on what the first left in the array), then another
    # shape and dtype (the array is cleared again), a greedy call (the arg-max cells) and the first shape once more
    plan = [(11, 16, 4, 32000, torch.float32, 1.0, 0.9), (12, 16, 4, 32000, torch.float32, 1.0, 0.9),
            (13, 5, 3, 6000, torch.bfloat16, 1.0, 0.0), (14, 16, 4, 32000, torch.float32, 0.0, 0.0),
            (11, 16, 4, 32000, torch.float32, 1.0, 0.9)]
    for step, (seed, B, gamma, V, dtype, T, top_p) in enumerate(plan):
        dl, tl, toks, r, noise = (t.to(dev) for t in _case(seed, B, gamma, V, dtype, top_p if T > 0 else 0.9))
        want = {k: v.clone() for k, v in ops.verify_fused(dl, tl, toks, r, noise, T, 0, top_p).items()}
        out = _outputs(B, gamma, dev)
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated(dev)
        rc = _call_static(lib, dl, tl, toks, r, noise, T, 0, top_p, out)
        assert rc == 0, (step, rc)
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated(dev) == before  # (torch's allocator was not involved either)
        _same(out, want, f"call {step}")
    # beyond the built-in array: a clear error, nothing launched, outputs untouched
    B, gamma = 40000, 8
    assert lib.duodec_verify_workspace_bytes(B, gamma) > lib.duodec_verify_static_max_bytes()
    small = _outputs(1, 1, dev)
    dl, tl, toks, r, noise = (t.to(dev) for t in _case(15, 1, 1, 64, torch.float32, 0.0))
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    rc = lib.duodec_verify_fused_static(
        p(dl), p(tl), 0, 0, 64, 0, 64, B, gamma, 64, 64, p(toks), 0, p(r), 0, p(noise), 0, 1.0, 0, 0.0, 1.0, 0, 0.0, 0,
        p(small["accept_mask"]), p(small["selected_target_p"]), p(small["selected_draft_p"]), p(small["accepted_len"]),
        p(small["token"]), p(small["all_accepted"]), None)
    assert rc == -4
    with pytest.raises(_lib.DuodecError, match="DUODEC_ESTATICMAX"):
        _lib.check(rc, "duodec_verify_fused_static")
    torch.cuda.synchronize()
    assert int(small["accepted_len"][0]) == -1
