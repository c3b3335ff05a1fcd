"""Host-buffer front end of the fused verification step.

The reference's callers hand the verification ops tensors that may live on
another device than the one that verifies (``prepare_verification_inputs``
moves the target rows to the draft device, /root/reference/src/decoding_ops.py:
88-100) and read Python ints back (``accepted_count``, ``n``).  This class is
that hop for logits that sit in (pinned) HOST memory: it splits the batch into
``nsplit`` groups of sequences, streams group i+1 over PCIe on a copy stream
while group i is being verified, and returns the small per-sequence results in
pinned host memory after ONE stream synchronisation.  All arithmetic is the
C-ABI kernel behind :func:`duodecoding_b200.ops.verify_fused`.
"""

from __future__ import annotations

from typing import Dict, Optional

import torch

from . import ops


class HostVerifier:
    """Verify ``[B, gamma(+1), V]`` host logits; device staging is allocated once."""

    def __init__(self, B: int, gamma: int, V: int, dtype: torch.dtype = torch.float32,
                 device: Optional[torch.device] = None, nsplit: int = 4):
        if not torch.cuda.is_available():
            raise RuntimeError("HostVerifier needs a CUDA device (duodecoding_b200 has no CPU path)")
        self.B, self.gamma, self.V, self.dtype = int(B), int(gamma), int(V), dtype
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        nsplit = max(1, min(int(nsplit), self.B))
        edges = [round(i * self.B / nsplit) for i in range(nsplit + 1)]
        self.groups = [(a, b) for a, b in zip(edges[:-1], edges[1:]) if b > a]
        d = self.device
        self.d_draft = torch.empty((B, gamma, V), dtype=dtype, device=d)
        self.d_target = torch.empty((B, gamma + 1, V), dtype=dtype, device=d)
        self.d_tokens = torch.empty((B, gamma), dtype=torch.int64, device=d)
        self.d_r = torch.empty((B, gamma), dtype=torch.float32, device=d)
        self.d_noise = torch.empty((B, V), dtype=torch.float32, device=d)
        self.d_out = {
            "accept_mask": torch.empty((B, gamma), dtype=torch.bool, device=d),
            "selected_target_p": torch.empty((B, gamma), dtype=torch.float32, device=d),
            "selected_draft_p": torch.empty((B, gamma), dtype=torch.float32, device=d),
            "accepted_len": torch.empty((B,), dtype=torch.int64, device=d),
            "token": torch.empty((B,), dtype=torch.int64, device=d),
            "all_accepted": torch.empty((B,), dtype=torch.bool, device=d),
        }
        self.h_out = {k: torch.empty(v.shape, dtype=v.dtype).pin_memory() for k, v in self.d_out.items()}
        self.copy_stream = torch.cuda.Stream(device=d)
        self.ready = [torch.cuda.Event() for _ in self.groups]
        self.h2d_bytes = 0
        self.d2h_bytes = sum(v.numel() * v.element_size() for v in self.h_out.values())
        self.kernel_launches = 0

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
        if noise is None:
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
        for k, v in self.d_out.items():
            self.h_out[k].copy_(v, non_blocking=True)
        main.synchronize()  # the caller reads Python ints next (src/decoding_ops.py:146)
        self.kernel_launches = launches
        return self.h_out
