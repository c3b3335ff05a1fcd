"""Host-buffer front end of the fused verification step.

The reference's callers hand the verification ops tensors that may live on
another device than the one that verifies (``prepare_verification_inputs``
moves the target rows to the draft device, /root/reference/src/decoding_ops.py:
88-100) and read Python ints back (``accepted_count``, ``n``).  This class is
that hop for logits that sit in (pinned) HOST memory.  The verification itself
takes ~50 us at the C2 shape while the logits need ~1.5 ms of PCIe time, so the
only thing that matters here is the copy: ONE ``cudaMemcpyAsync`` per step at
full link speed.  ``host_set()`` hands out pinned views (draft, target, tokens,
r, noise) over ONE pinned arena that mirrors the device arena; the producer of
the logits writes into them and ``verify_set`` moves the arena with a single
copy (small or strided copies reach half the link speed or less:
``tools/h2d_probe.py``).  ``verify`` accepts arbitrary host tensors and copies
tensor by tensor (whole contiguous tensors; a ``nsplit`` > 1 pipelines groups
of sequences, which only pays for batches whose verification time is
comparable to their copy time).  Results come back in pinned host memory after
ONE device-to-host copy and ONE stream synchronisation.  All arithmetic is the
C-ABI kernel behind :func:`duodecoding_b200.ops.verify_fused`.
"""

from __future__ import annotations

from typing import Dict, Optional

import torch

from . import ops


def _align(n: int, a: int = 256) -> int:
    return (n + a - 1) // a * a


class HostVerifier:
    """Verify ``[B, gamma(+1), V]`` host logits; device and host staging are allocated once."""

    _OUT = (("accept_mask", torch.bool, "g"), ("selected_target_p", torch.float32, "g"),
            ("selected_draft_p", torch.float32, "g"), ("accepted_len", torch.int64, "1"),
            ("token", torch.int64, "1"), ("all_accepted", torch.bool, "1"))

    def __init__(self, B: int, gamma: int, V: int, dtype: torch.dtype = torch.float32,
                 device: Optional[torch.device] = None, nsplit: int = 1):
        if not torch.cuda.is_available():
            raise RuntimeError("HostVerifier needs a CUDA device (duodecoding_b200 has no CPU path)")
        self.B, self.gamma, self.V, self.dtype = int(B), int(gamma), int(V), dtype
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        nsplit = max(1, min(int(nsplit), self.B))
        edges = [round(i * self.B / nsplit) for i in range(nsplit + 1)]
        self.groups = [(a, b) for a, b in zip(edges[:-1], edges[1:]) if b > a]
        d = self.device
        elt = torch.empty((), dtype=dtype).element_size()
        # ---- one input arena: draft | target | tokens | r | noise (the optional noise last: the copy stops before it)
        self._fields, off = {}, 0
        for name, shape, dt, size in (("draft", (B, gamma, V), dtype, elt), ("target", (B, gamma + 1, V), dtype, elt),
                                      ("tokens", (B, gamma), torch.int64, 8), ("r", (B, gamma), torch.float32, 4),
                                      ("noise", (B, V), torch.float32, 4)):
            n = size
            for x in shape:
                n *= x
            self._fields[name] = (off, n, shape, dt)
            if name == "r":
                self._in_bytes = off + n
            off = _align(off + n)
        self._arena_bytes = off
        self.d_arena = torch.empty(self._arena_bytes, dtype=torch.uint8, device=d)
        v = self._views(self.d_arena)
        self.d_draft, self.d_target, self.d_tokens, self.d_r, self.d_noise = (
            v["draft"], v["target"], v["tokens"], v["r"], v["noise"])
        # ---- one output arena
        self._ofields, off = {}, 0
        for name, dt, kind in self._OUT:
            shape = (B, gamma) if kind == "g" else (B,)
            n = torch.empty((), dtype=dt).element_size() * B * (gamma if kind == "g" else 1)
            self._ofields[name] = (off, n, shape, dt)
            off = _align(off + n, 16)
        self.d_oarena = torch.empty(max(off, 16), dtype=torch.uint8, device=d)
        self.h_oarena = torch.empty(max(off, 16), dtype=torch.uint8).pin_memory()
        self.d_out = {k: self.d_oarena[o:o + n].view(dt).view(shape) for k, (o, n, shape, dt) in self._ofields.items()}
        self.h_out = {k: self.h_oarena[o:o + n].view(dt).view(shape) for k, (o, n, shape, dt) in self._ofields.items()}
        self.copy_stream = torch.cuda.Stream(device=d)
        self.ready = [torch.cuda.Event() for _ in self.groups]
        self.h2d_bytes = 0
        self.d2h_bytes = sum(n for _, n, _, _ in self._ofields.values())
        self.kernel_launches = 0

    def _views(self, arena: torch.Tensor) -> Dict[str, torch.Tensor]:
        return {k: arena[o:o + n].view(dt).view(shape) for k, (o, n, shape, dt) in self._fields.items()}

    def host_set(self) -> Dict[str, torch.Tensor]:
        """Pinned host views ``draft [B, gamma, V]``, ``target [B, gamma+1, V]``, ``tokens``, ``r``, ``noise`` over one
        pinned arena laid out like the device arena: fill them, then :meth:`verify_set`.  Several sets may be kept
        (double buffering on the producer's side)."""
        arena = torch.empty(self._arena_bytes, dtype=torch.uint8).pin_memory()
        hs = self._views(arena)
        hs["_arena"] = arena
        return hs

    def _finish(self, noise_given: bool, temperature: float, top_k: int, top_p: float) -> Dict[str, torch.Tensor]:
        main = torch.cuda.current_stream(self.device)
        if not noise_given:
            # the Exp(1) draws torch.multinomial would make on the device (src/utils.py:747)
            self.d_noise.exponential_(1.0)
        launches = 0
        for i, (a, b) in enumerate(self.groups):
            main.wait_event(self.ready[i])
            ops.verify_fused(
                self.d_draft[a:b], self.d_target[a:b], self.d_tokens[a:b], self.d_r[a:b], self.d_noise[a:b],
                temperature, top_k, top_p, out={k: v[a:b] for k, v in self.d_out.items()},
            )
            launches += 1
        self.h_oarena.copy_(self.d_oarena, non_blocking=True)
        main.synchronize()  # the caller reads Python ints next (src/decoding_ops.py:146)
        self.kernel_launches = launches
        return self.h_out

    def verify_set(self, hs: Dict[str, torch.Tensor], with_noise: bool = False, temperature: float = 1.0,
                   top_k: int = 0, top_p: float = 0.0) -> Dict[str, torch.Tensor]:
        """One step from a :meth:`host_set`: ONE host-to-device copy of the arena (the noise part only when
        ``with_noise``), the fused verification, ONE device-to-host copy of the results."""
        main = torch.cuda.current_stream(self.device)
        self.copy_stream.wait_stream(main)  # the device arena of the previous step is free
        n = self._arena_bytes if with_noise else self._in_bytes
        with torch.cuda.stream(self.copy_stream):
            self.d_arena[:n].copy_(hs["_arena"][:n], non_blocking=True)
            for ev in self.ready:
                ev.record(self.copy_stream)
        self.h2d_bytes = (self._fields["noise"][1] if with_noise else 0) + sum(
            self._fields[k][1] for k in ("draft", "target", "tokens", "r"))
        return self._finish(with_noise, temperature, top_k, top_p)

    def verify(self, draft: torch.Tensor, target: torch.Tensor, tokens: torch.Tensor, r: torch.Tensor,
               noise: Optional[torch.Tensor] = None, temperature: float = 1.0, top_k: int = 0,
               top_p: float = 0.0) -> Dict[str, torch.Tensor]:
        """Host tensors in (pinned for full PCIe speed), pinned host results out."""
        g = self.gamma
        main = torch.cuda.current_stream(self.device)
        self.copy_stream.wait_stream(main)  # staging buffers of the previous step are free
        nbytes = 0
        with torch.cuda.stream(self.copy_stream):
            self.d_tokens.copy_(tokens, non_blocking=True)
            self.d_r.copy_(r, non_blocking=True)
            nbytes += tokens.numel() * 8 + r.numel() * 4
            for i, (a, b) in enumerate(self.groups):
                self.d_draft[a:b].copy_(draft[a:b, :g], non_blocking=True)
                self.d_target[a:b].copy_(target[a:b, : g + 1], non_blocking=True)
                nbytes += (b - a) * (2 * g + 1) * self.V * self.d_draft.element_size()
                if noise is not None:
                    self.d_noise[a:b].copy_(noise[a:b], non_blocking=True)
                    nbytes += (b - a) * self.V * 4
                self.ready[i].record(self.copy_stream)
        self.h2d_bytes = nbytes
        return self._finish(noise is not None, temperature, top_k, top_p)
