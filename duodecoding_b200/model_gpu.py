"""Stateful model cache of the verification path (mirror of /root/reference/src/model_gpu.py).

``KVCacheModel`` keeps the reference's surface (ctor, ``vocab_size``, ``prob_history``,
``logits_history``, ``generate``, ``generate_with_rebuilt_topk``, ``rollback`` ...), with

* logits -> probabilities by the ``duodec_norm_logits`` kernel, written straight into the history
  buffer (the reference makes 3-12 elementwise passes and two extra V-sized copies per row,
  src/model_gpu.py:186-203, :246-258);
* rollback = truncation in place (src/model_gpu.py:362-383): the transformer cache is cropped
  through its own ``crop`` (views only) and the per-sequence length vector ``seq_lens`` - the device
  side record a static cache reads - is clamped by the ``duodec_kv_truncate`` kernel; no cache byte
  moves.  ``StaticKVCache`` below is that static cache as a stand-alone ``CacheLike``.

The model forward itself (``self._model(...)``) is below the hot path and stays the caller's.
"""

from __future__ import annotations

from typing import Iterator, List, Optional, Tuple

import torch

from . import ops
from .utils import norm_logits, rebuild_topk_uniform_probs, sample  # noqa: F401  (patch points)


class StaticKVCache:
    """Pre-allocated K/V buffers with per-sequence lengths: the reference's ``CacheLike`` protocol
    (``crop`` / ``get_seq_length`` / ``__iter__``, src/model_gpu.py:23-26) where ``crop`` is an integer
    write on the device (``duodec_kv_truncate``) instead of slicing or re-concatenating tensors."""

    def __init__(self, num_layers: int, batch: int, heads: int, max_len: int, head_dim: int,
                 dtype: torch.dtype = torch.bfloat16, device: Optional[torch.device] = None):
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        shape = (batch, heads, max_len, head_dim)
        self.keys = [torch.empty(shape, dtype=dtype, device=dev) for _ in range(num_layers)]
        self.values = [torch.empty(shape, dtype=dtype, device=dev) for _ in range(num_layers)]
        self.seq_lens = torch.zeros((batch,), dtype=torch.int32, device=dev)  # what attention kernels read
        self.max_len = max_len
        self._len = 0  # host mirror of the (uniform) length, for callers that need a Python int

    def append(self, layer: int, k: torch.Tensor, v: torch.Tensor) -> None:
        """Write ``[B, H, S, D]`` new entries of one layer at the current length (in place)."""
        s = k.shape[2]
        if self._len + s > self.max_len:
            raise ValueError("StaticKVCache overflow")
        self.keys[layer][:, :, self._len:self._len + s].copy_(k)
        self.values[layer][:, :, self._len:self._len + s].copy_(v)
        if layer == len(self.keys) - 1:
            self._len += s
            self.seq_lens.fill_(self._len)

    def crop(self, end_pos) -> None:
        """Truncate to ``end_pos`` (an int, or an int64 tensor with one entry per sequence)."""
        ops.kv_truncate(self.seq_lens, end_pos)
        if not isinstance(end_pos, torch.Tensor):
            self._len = min(self._len, max(int(end_pos), 0))

    def get_seq_length(self, layer_idx: int = 0) -> int:
        return self._len

    def __iter__(self) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        for k, v in zip(self.keys, self.values):
            yield k[:, :, :self._len], v[:, :, :self._len]

    def __len__(self) -> int:
        return len(self.keys)


class KVCacheModel:
    """src/model_gpu.py:52-422, same constructor and attributes."""

    def __init__(self, model, temperature: float = 1, top_k: int = 0, top_p: float = 0,
                 return_hidden_states: bool = False, max_length: int = 16384) -> None:
        self._model = model
        self._past_key_values = None
        self._temperature, self._top_k, self._top_p = temperature, top_k, top_p
        self.max_length = max_length
        self.hidden_states = None
        self.embedding_vocab_size = int(model.get_input_embeddings().weight.shape[0])
        cfg = model.config
        if hasattr(cfg, "vocab_size"):
            self.vocab_size = int(cfg.vocab_size)
        elif hasattr(cfg, "text_config") and hasattr(cfg.text_config, "vocab_size"):
            self.vocab_size = int(cfg.text_config.vocab_size)
        else:
            raise AttributeError("Vocab size not found in model config")
        self._prob_buffer: Optional[torch.Tensor] = None
        self._logits_buffer: Optional[torch.Tensor] = None
        self._current_seq_len = 0
        self.seq_lens: Optional[torch.Tensor] = None  # int32 [B] on the device, clamped by rollback

    # ---- history views (writable: callers assign prob_history[:, n, :V] = ..., src/engine.py:948-954)
    @property
    def _prob_history(self) -> Optional[torch.Tensor]:
        return None if self._prob_buffer is None else self._prob_buffer[:, :self._current_seq_len, :]

    @_prob_history.setter
    def _prob_history(self, value) -> None:
        pass

    @property
    def logits_history(self) -> Optional[torch.Tensor]:
        return None if self._logits_buffer is None else self._logits_buffer[:, :self._current_seq_len, :]

    @logits_history.setter
    def logits_history(self, value) -> None:
        pass

    @property
    def prob_history(self) -> torch.Tensor:
        hist = self._prob_history
        if hist is None:
            raise ValueError("Probability history buffer is not initialized")
        return hist

    def _ensure_buffer_size(self, batch_size: int, seq_len: int, device, dtype) -> None:
        """Geometric growth of the two [B, L, V] history buffers (src/model_gpu.py:116-153)."""
        if self._prob_buffer is None:
            self.max_length = max(2048, seq_len + 1024)
            shape = (batch_size, self.max_length, self.vocab_size)
            self._prob_buffer = torch.empty(shape, device=device, dtype=dtype)
            self._logits_buffer = torch.empty(shape, device=device, dtype=dtype)
            self.seq_lens = torch.zeros((batch_size,), dtype=torch.int32, device=device)
            return
        if seq_len > self.max_length:
            new_len = max(self.max_length * 2, seq_len + 1024)
            for name in ("_prob_buffer", "_logits_buffer"):
                old = getattr(self, name)
                new = torch.empty((batch_size, new_len, self.vocab_size), device=device, dtype=dtype)
                new[:, :self._current_seq_len].copy_(old[:, :self._current_seq_len])
                setattr(self, name, new)
            self.max_length = new_len

    def _validate_input_ids(self, input_ids: torch.Tensor) -> None:
        """src/model_gpu.py:267-282: ValueError for ids outside the embedding table."""
        if input_ids.numel() == 0:
            return
        lo, hi = (int(v) for v in torch.aminmax(input_ids))
        if lo < 0 or hi >= self.embedding_vocab_size:
            name = getattr(getattr(self._model, "config", None), "_name_or_path", "unknown")
            raise ValueError(
                f"Input token id out of embedding range for model {name}: min={lo}, max={hi}, "
                f"embedding_vocab_size={self.embedding_vocab_size}, configured_vocab_size={self.vocab_size}")

    def _store_rows(self, logits: torch.Tensor, start: int) -> torch.Tensor:
        """Logits [B, S, >=V] -> both history buffers at rows [start, start+S); returns the last probs."""
        B, S = logits.shape[0], logits.shape[1]
        sliced = logits[..., :self.vocab_size]
        self._logits_buffer[:, start:start + S, :] = sliced
        for b in range(B):  # the kernel writes the probabilities straight into the history rows
            ops.norm_logits(sliced[b], float(self._temperature), int(self._top_k or 0), float(self._top_p or 0.0),
                            out=self._prob_buffer[b, start:start + S, :])
        self._current_seq_len = start + S
        self.seq_lens.fill_(self._current_seq_len)
        return self._prob_buffer[:, start + S - 1, :]

    @torch.no_grad()
    def _prefill(self, input_ids: torch.Tensor) -> torch.Tensor:
        """src/model_gpu.py:172-208: whole prompt through the model; [B, V] next-token probabilities."""
        self._validate_input_ids(input_ids)
        outputs = self._model(input_ids)
        logits = outputs.logits
        if logits is None:
            raise RuntimeError("Model returned logits=None in prefill")
        self._ensure_buffer_size(input_ids.shape[0], input_ids.shape[1], logits.device, logits.dtype)
        probs = self._store_rows(logits, 0)
        self._past_key_values = outputs.past_key_values
        self.hidden_states = outputs.hidden_states
        return probs

    @torch.no_grad()
    def _decode_step(self, last_input_id: torch.Tensor) -> torch.Tensor:
        """src/model_gpu.py:210-265: one cached step (or a short cached suffix)."""
        if last_input_id.dtype != torch.long:
            last_input_id = last_input_id.to(torch.long)
        self._validate_input_ids(last_input_id)
        if last_input_id.shape[1] == 0:
            if self._current_seq_len <= 0:
                raise RuntimeError("No new input provided for decode step and no cache available")
            return self.prob_history[:, self._current_seq_len - 1, :]
        if self._past_key_values is None:
            raise RuntimeError("Decode step called before cache initialization")
        outputs = self._model(last_input_id, past_key_values=self._past_key_values, use_cache=True)
        logits = outputs.logits
        if logits is None:
            raise RuntimeError("Model returned logits=None in decode step")
        if outputs.past_key_values is None:
            raise RuntimeError("Model returned past_key_values=None in decode step")
        end_pos = self._current_seq_len + last_input_id.shape[1]
        self._ensure_buffer_size(last_input_id.shape[0], end_pos, logits.device, logits.dtype)
        probs = self._store_rows(logits, self._current_seq_len)
        self._past_key_values = outputs.past_key_values
        self.hidden_states = outputs.hidden_states
        return probs

    def _forward_with_kvcache(self, input_ids: torch.Tensor) -> torch.Tensor:
        """src/model_gpu.py:284-300."""
        if input_ids.dtype != torch.long:
            input_ids = input_ids.to(torch.long)
        if self._past_key_values is None:
            return self._prefill(input_ids)
        return self._decode_step(input_ids[:, self.current_length:])

    @property
    def last_hidden_state(self) -> torch.Tensor:
        if self.hidden_states is None:
            raise ValueError("hidden_states is None")
        return self.hidden_states[-1]

    def _generate_with_kvcache(self, prefix: torch.Tensor, gamma: int) -> torch.Tensor:
        """src/model_gpu.py:308-322: gamma x (forward -> sample)."""
        x = prefix if prefix.dtype == torch.long else prefix.to(torch.long)
        for _ in range(gamma):
            x = torch.cat((x, sample(self._forward_with_kvcache(x)).to(torch.long)), dim=1)
        return x

    def generate_with_rebuilt_topk(self, input: torch.Tensor, gamma: int,
                                   proposal_top_k: Optional[int]) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
        """src/model_gpu.py:324-356: gamma x (forward -> top-k rebuild -> sample); the rebuilt rows
        are written into one pre-sized [B, gamma, V] tensor instead of a list + cat."""
        x = input if input.dtype == torch.long else input.to(torch.long)
        rebuilt: Optional[torch.Tensor] = None
        for j in range(gamma):
            q = rebuild_topk_uniform_probs(self._forward_with_kvcache(x), proposal_top_k)
            if rebuilt is None:
                rebuilt = torch.empty((q.shape[0], gamma, q.shape[-1]), dtype=q.dtype, device=q.device)
            rebuilt[:, j, :] = q
            x = torch.cat((x, sample(q).to(torch.long)), dim=1)
        return x, rebuilt

    @torch.no_grad()
    def generate(self, input: torch.Tensor, gamma: int) -> torch.Tensor:
        return self._generate_with_kvcache(input, gamma)

    @torch.no_grad()
    def rollback(self, end_pos: int) -> None:
        """src/model_gpu.py:362-383: truncate in place; lengths only, no cache bytes move."""
        past = self._past_key_values
        if past is None:
            return
        if hasattr(past, "crop"):
            past.crop(end_pos)
        else:  # legacy tuple of (k, v): views, like the reference
            self._past_key_values = tuple((k[:, :, :end_pos, :], v[:, :, :end_pos, :]) for k, v in past)
        self._current_seq_len = min(end_pos, self.current_length)
        if self.seq_lens is not None:
            ops.kv_truncate(self.seq_lens, int(end_pos))

    @property
    def device(self) -> torch.device:
        return next(iter(self._model.parameters())).device

    @property
    def current_length(self) -> int:
        past = self._past_key_values
        if past is None:
            return 0
        if hasattr(past, "get_seq_length"):
            return int(past.get_seq_length())
        return int(past[0][0].shape[2])

    def debug_state(self) -> dict:
        have = 0 if self._prob_buffer is None else self._current_seq_len
        return {"current_length": self.current_length, "tracked_seq_len": self._current_seq_len,
                "prob_history_len": have, "logits_history_len": have, "max_length": self.max_length}

    def debug_row_sums(self, start: int, end: int) -> List[float]:
        if self._prob_buffer is None or self._current_seq_len <= 0:
            return []
        start = max(0, min(start, self._current_seq_len))
        end = max(start, min(end, self._current_seq_len))
        return self._prob_buffer[0, start:end, :].float().sum(dim=-1).tolist() if end > start else []

    def __len__(self) -> int:
        return self.current_length
