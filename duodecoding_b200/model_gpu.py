"""Stateful model cache of the verification path (mirror of /root/reference/src/model_gpu.py).

``KVCacheModel`` keeps the reference's surface (ctor, ``vocab_size``, ``prob_history``,
``logits_history``, ``generate``, ``generate_with_rebuilt_topk``, ``rollback`` ...).  What changed
underneath:

* **logits are the history; probabilities are lazy.**  A forward stores the new logit rows (one
  strided copy).  ``softmax(filter(logits / T))`` rows are only materialised - by ONE batched
  ``duodec_norm_logits`` launch over all pending rows, written straight into the probability buffer -
  when somebody reads ``prob_history``.  The default decode loop never does: the draft's next token is
  drawn from the logits row by the fused kernel (``verify_fused`` with gamma = 0 is "sample from
  softmax(filter(logits / T))"), and verification (``decoding_ops``) runs on the logits rows.  The
  reference makes 3-12 elementwise passes plus two V-sized copies per row (src/model_gpu.py:186-203,
  :246-258).
* **history window** (``history_window=W``, off by default): verification only reads rows
  ``[P-1, P+gamma)`` (src/decoding_ops.py:88-100), so only the last W rows are kept, in a sliding
  buffer of 2W rows; ``prob_history`` / ``logits_history`` then return a row-window proxy that serves
  the slicing the reference's callers do and raises ``IndexError`` for rows that left the window.
  With it a cache of B=256, V=151936 needs ~1.2 GB instead of 0.6 GB per sequence and model.
* **KV rollback is an integer write.**  The transformer cache handed to the model is an
  ``InPlaceKVCache``: K/V live in pre-allocated buffers, ``update`` writes the new positions in place
  and returns views, ``crop`` only sets the length (the reference's ``DynamicCache`` re-``cat``s all
  K/V on every forward and slices on crop, src/model_gpu.py:362-383).

The model forward itself (``self._model(...)``) is below the hot path and stays the caller's.
"""

from __future__ import annotations

from typing import Iterator, List, Optional, Tuple

import torch
from transformers.cache_utils import Cache, DynamicLayer

from . import ops
from .utils import norm_logits, rebuild_topk_uniform_probs, sample  # noqa: F401  (patch points)

_ORIG_SAMPLE = sample


class InPlaceLayer(DynamicLayer):
    """One layer of K/V in pre-allocated ``[B, H, cap, D]`` buffers.  ``update`` copies the new
    positions to ``[len, len+S)`` and returns views of ``[:len+S]``; ``crop`` sets the length.  No
    ``torch.cat``, no cache bytes move on rollback."""

    initial_capacity = 512

    def lazy_initialization(self, key_states: torch.Tensor, value_states: torch.Tensor) -> None:
        self.dtype, self.device = key_states.dtype, key_states.device
        b, h, s, d = key_states.shape
        cap = max(self.initial_capacity, 2 * s)
        self._kbuf = torch.empty((b, h, cap, d), dtype=self.dtype, device=self.device)
        self._vbuf = torch.empty((b, h, cap, value_states.shape[-1]), dtype=value_states.dtype, device=self.device)
        self._len = 0
        self.keys, self.values = self._kbuf[:, :, :0], self._vbuf[:, :, :0]
        self.is_initialized = True

    def _grow(self, need: int) -> None:
        cap = max(2 * self._kbuf.shape[2], need)
        for name in ("_kbuf", "_vbuf"):
            old = getattr(self, name)
            new = torch.empty((old.shape[0], old.shape[1], cap, old.shape[3]), dtype=old.dtype, device=old.device)
            new[:, :, :self._len].copy_(old[:, :, :self._len])
            setattr(self, name, new)

    def update(self, key_states: torch.Tensor, value_states: torch.Tensor, *args, **kwargs):
        if not self.is_initialized:
            self.lazy_initialization(key_states, value_states)
        end = self._len + key_states.shape[-2]
        if end > self._kbuf.shape[2]:
            self._grow(end)
        self._kbuf[:, :, self._len:end].copy_(key_states)
        self._vbuf[:, :, self._len:end].copy_(value_states)
        self._len = end
        self.keys, self.values = self._kbuf[:, :, :end], self._vbuf[:, :, :end]
        return self.keys, self.values

    def get_seq_length(self) -> int:
        return self._len if self.is_initialized else 0

    def crop(self, max_length: int) -> None:
        if not self.is_initialized:
            return
        if max_length < 0:
            max_length = self._len - abs(max_length)
        if self._len <= max_length:
            return
        self._len = max(int(max_length), 0)
        self.keys, self.values = self._kbuf[:, :, :self._len], self._vbuf[:, :, :self._len]


class InPlaceKVCache(Cache):
    """The reference's ``CacheLike`` protocol (``crop`` / ``get_seq_length`` / ``__iter__``,
    src/model_gpu.py:23-26) as a transformers ``Cache`` whose layers update in place.  ``seq_lens``
    (int32 ``[B]`` on the device, clamped by the ``duodec_kv_truncate`` kernel on ``crop``) is the
    ``cache_seqlens`` operand of attention kernels that take per-sequence lengths; ``crop`` also accepts
    an int64 ``[B]`` tensor of per-sequence end positions for it."""

    def __init__(self) -> None:
        super().__init__(layer_class_to_replicate=InPlaceLayer)
        self.seq_lens: Optional[torch.Tensor] = None

    def update(self, key_states, value_states, layer_idx, *args, **kwargs):
        out = super().update(key_states, value_states, layer_idx, *args, **kwargs)
        if layer_idx == 0 and key_states.is_cuda:
            if self.seq_lens is None:
                self.seq_lens = torch.zeros((key_states.shape[0],), dtype=torch.int32, device=key_states.device)
            self.seq_lens.fill_(self.layers[0].get_seq_length())
        return out

    def crop(self, max_length) -> None:
        if isinstance(max_length, torch.Tensor):  # ragged: device lengths only, the uniform length is the largest
            if self.seq_lens is not None:
                ops.kv_truncate(self.seq_lens, max_length)
            return
        for layer in self.layers:
            layer.crop(int(max_length))
        if self.seq_lens is not None:
            ops.kv_truncate(self.seq_lens, int(max_length) if max_length >= 0 else self.get_seq_length())

    def get_seq_length(self, layer_idx: int = 0) -> int:
        return self.layers[layer_idx].get_seq_length() if layer_idx < len(self.layers) else 0

    def __iter__(self) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        for layer in self.layers:
            yield layer.keys, layer.values

    def __len__(self) -> int:
        return len(self.layers)


StaticKVCache = InPlaceKVCache  # round-1 name


class RowWindow:
    """History proxy of a windowed ``KVCacheModel``: behaves like the ``[B, L, V]`` history tensor for
    the accesses the reference's callers make - ``.shape`` / ``.device`` / ``.dtype``,
    ``h[:, a:b, :v]``, ``h[:, n, :v]``, ``h[:, -1, :v]`` and the same on the left of an assignment -
    with absolute row numbers; rows that left the window raise ``IndexError``."""

    def __init__(self, owner: "KVCacheModel", kind: str) -> None:
        self._o, self._kind = owner, kind

    def _buf(self) -> torch.Tensor:
        return self._o._window_tensor(self._kind)

    @property
    def shape(self) -> torch.Size:
        b = self._o._logits_buffer
        return torch.Size((b.shape[0], self._o._current_seq_len, b.shape[2]))

    def size(self, dim: Optional[int] = None):
        return self.shape if dim is None else self.shape[dim]

    @property
    def device(self) -> torch.device:
        return self._o._logits_buffer.device

    @property
    def dtype(self) -> torch.dtype:
        return self._o._logits_buffer.dtype

    def dim(self) -> int:
        return 3

    def _rows(self, key):
        if not isinstance(key, tuple):
            key = (key,)
        key = key + (slice(None),) * (3 - len(key))
        bsel, rsel, vsel = key
        L, base = self._o._current_seq_len, self._o._win_base
        if isinstance(rsel, slice):
            a, b, step = rsel.indices(L)
            if step != 1:
                raise IndexError("row window: unit row steps only")
            if b <= a:
                return bsel, slice(0, 0), vsel
            if a < base:
                raise IndexError(f"history rows [{a}, {b}) requested, rows below {base} left the window "
                                 f"(history_window={self._o.history_window})")
            return bsel, slice(a - base, b - base), vsel
        r = int(rsel)
        r = r + L if r < 0 else r
        if not 0 <= r < L:
            raise IndexError(f"history row {rsel} out of range ({L} rows)")
        if r < base:
            raise IndexError(f"history row {r} left the window (rows >= {base} are kept)")
        return bsel, r - base, vsel

    def __getitem__(self, key) -> torch.Tensor:
        return self._buf()[self._rows(key)]

    def __setitem__(self, key, value) -> None:
        self._buf()[self._rows(key)] = value


class KVCacheModel:
    """src/model_gpu.py:52-422, same constructor and attributes (+ keyword-only ``history_window``)."""

    def __init__(self, model, temperature: float = 1, top_k: int = 0, top_p: float = 0,
                 return_hidden_states: bool = False, max_length: int = 16384, *,
                 history_window: Optional[int] = None) -> None:
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
        self.history_window = None if not history_window else max(int(history_window), 2)
        self._prob_buffer: Optional[torch.Tensor] = None    # allocated by the first prob_history read
        self._logits_buffer: Optional[torch.Tensor] = None  # [B, max_length, V], or [B, 2W, V] with a window
        self._current_seq_len = 0
        self._win_base = 0    # absolute row of buffer row 0 (always 0 without a window)
        self._prob_done = 0   # rows below this absolute index have their probabilities in _prob_buffer
        self._fused_token = None  # (n, token) left by a fused verify for finalize_verification (decoding_ops)

    # ---- storage -------------------------------------------------------------------------------
    def _cap_rows(self) -> int:
        return 0 if self._logits_buffer is None else self._logits_buffer.shape[1]

    def _ensure_buffer_size(self, batch_size: int, seq_len: int, device, dtype) -> None:
        """Room for rows [.., seq_len): geometric growth of the full buffers (src/model_gpu.py:116-153) or a
        slide of the window (the last W rows move to the front; one copy every W rows)."""
        W = self.history_window
        if self._logits_buffer is None:
            if W is None:
                self.max_length = max(2048, seq_len + 1024)
                rows = self.max_length
            else:
                rows = 2 * W
            self._logits_buffer = torch.empty((batch_size, rows, self.vocab_size), device=device, dtype=dtype)
            return
        if W is None:
            if seq_len > self.max_length:
                new_len = max(self.max_length * 2, seq_len + 1024)
                for name in ("_prob_buffer", "_logits_buffer"):
                    old = getattr(self, name)
                    if old is None:
                        continue
                    new = torch.empty((batch_size, new_len, self.vocab_size), device=device, dtype=old.dtype)
                    new[:, :self._current_seq_len].copy_(old[:, :self._current_seq_len])
                    setattr(self, name, new)
                self.max_length = new_len
            return
        if seq_len - self._win_base > 2 * W:  # slide: keep the last W rows
            keep_from = max(self._current_seq_len - W, self._win_base)
            n = self._current_seq_len - keep_from
            for name in ("_prob_buffer", "_logits_buffer"):
                buf = getattr(self, name)
                if buf is not None and n > 0:
                    buf[:, :n].copy_(buf[:, keep_from - self._win_base:keep_from - self._win_base + n].clone())
            self._win_base = keep_from
            self._prob_done = max(self._prob_done, self._win_base)

    def _materialize_probs(self) -> None:
        """Probabilities of every stored row that does not have them yet: one batched kernel launch,
        logits rows in, probability rows out, nothing else V-sized."""
        if self._logits_buffer is None:
            return
        if self._prob_buffer is None:
            self._prob_buffer = torch.empty_like(self._logits_buffer)
        lo, hi = max(self._prob_done, self._win_base), self._current_seq_len
        if hi > lo:
            a, b = lo - self._win_base, hi - self._win_base
            ops.norm_logits(self._logits_buffer[:, a:b], float(self._temperature), int(self._top_k or 0),
                            float(self._top_p or 0.0), out=self._prob_buffer[:, a:b])
        self._prob_done = max(self._prob_done, hi)

    def _window_tensor(self, kind: str) -> torch.Tensor:
        if kind == "prob":
            self._materialize_probs()
            return self._prob_buffer[:, :self._current_seq_len - self._win_base]
        return self._logits_buffer[:, :self._current_seq_len - self._win_base]

    # ---- history views (writable: callers assign prob_history[:, n, :V] = ..., src/engine.py:948-954)
    @property
    def _prob_history(self):
        if self._logits_buffer is None:
            return None
        self._fused_token = None  # whoever reads (or edits) probabilities gets the probability-space flow
        if self.history_window is not None:
            return RowWindow(self, "prob")
        self._materialize_probs()
        return self._prob_buffer[:, :self._current_seq_len, :]

    @_prob_history.setter
    def _prob_history(self, value) -> None:
        pass

    @property
    def logits_history(self):
        if self._logits_buffer is None:
            return None
        if self.history_window is not None:
            return RowWindow(self, "logits")
        return self._logits_buffer[:, :self._current_seq_len, :]

    @logits_history.setter
    def logits_history(self, value) -> None:
        pass

    @property
    def prob_history(self):
        hist = self._prob_history
        if hist is None:
            raise ValueError("Probability history buffer is not initialized")
        return hist

    def logits_rows(self, start: int, end: int) -> Optional[torch.Tensor]:
        """Stored logit rows [start, end) as a plain strided view (None when they are not all kept):
        what the fused verification reads."""
        if self._logits_buffer is None or start < self._win_base or end > self._current_seq_len or end < start:
            return None
        return self._logits_buffer[:, start - self._win_base:end - self._win_base, :]

    def _last_probs(self) -> torch.Tensor:
        """[B, V] probabilities of the newest row (what _prefill / _decode_step return)."""
        return self.prob_history[:, self._current_seq_len - 1, :]

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

    def _store_rows(self, logits: torch.Tensor, start: int) -> None:
        """Logits [B, S, >=V] -> history rows [start, start+S) (one strided copy; no probabilities yet)."""
        S = logits.shape[1]
        if self.history_window is not None and S > 2 * self.history_window:  # a long prompt: only its tail is kept
            drop = S - self.history_window
            logits, start, S = logits[:, drop:], start + drop, self.history_window
            self._win_base, self._current_seq_len = start, start
        self._ensure_buffer_size(logits.shape[0], start + S, logits.device, logits.dtype)
        a = start - self._win_base
        self._logits_buffer[:, a:a + S, :].copy_(logits[..., :self.vocab_size])
        self._current_seq_len = start + S
        self._prob_done = min(self._prob_done, start)
        self._fused_token = None

    def _forward_logits(self, input_ids: torch.Tensor) -> None:
        """Prefill or cached step: run the model, store the new logit rows (src/model_gpu.py:172-265)."""
        if input_ids.dtype != torch.long:
            input_ids = input_ids.to(torch.long)
        if self._past_key_values is None:
            self._validate_input_ids(input_ids)
            outputs = self._model(input_ids, past_key_values=InPlaceKVCache(), use_cache=True)
            if outputs.logits is None:
                raise RuntimeError("Model returned logits=None in prefill")
            start = 0
        else:
            new_ids = input_ids[:, self.current_length:]
            self._validate_input_ids(new_ids)
            if new_ids.shape[1] == 0:
                if self._current_seq_len <= 0:
                    raise RuntimeError("No new input provided for decode step and no cache available")
                return
            outputs = self._model(new_ids, past_key_values=self._past_key_values, use_cache=True)
            if outputs.logits is None:
                raise RuntimeError("Model returned logits=None in decode step")
            if outputs.past_key_values is None:
                raise RuntimeError("Model returned past_key_values=None in decode step")
            start = self._current_seq_len
        self._store_rows(outputs.logits, start)
        self._past_key_values = outputs.past_key_values
        self.hidden_states = outputs.hidden_states

    @torch.no_grad()
    def _prefill(self, input_ids: torch.Tensor) -> torch.Tensor:
        """src/model_gpu.py:172-208: whole prompt through the model; [B, V] next-token probabilities."""
        self._past_key_values = None
        self._forward_logits(input_ids)
        return self._last_probs()

    @torch.no_grad()
    def _decode_step(self, last_input_id: torch.Tensor) -> torch.Tensor:
        """src/model_gpu.py:210-265: one cached step (or a short cached suffix)."""
        if self._past_key_values is None and last_input_id.shape[1] > 0:
            raise RuntimeError("Decode step called before cache initialization")
        if last_input_id.dtype != torch.long:
            last_input_id = last_input_id.to(torch.long)
        if last_input_id.shape[1] == 0:
            if self._current_seq_len <= 0:
                raise RuntimeError("No new input provided for decode step and no cache available")
            return self._last_probs()
        self._validate_input_ids(last_input_id)
        outputs = self._model(last_input_id, past_key_values=self._past_key_values, use_cache=True)
        if outputs.logits is None:
            raise RuntimeError("Model returned logits=None in decode step")
        if outputs.past_key_values is None:
            raise RuntimeError("Model returned past_key_values=None in decode step")
        self._store_rows(outputs.logits, self._current_seq_len)
        self._past_key_values = outputs.past_key_values
        self.hidden_states = outputs.hidden_states
        return self._last_probs()

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

    def _sample_last_row(self) -> torch.Tensor:
        """Next token [B, 1] drawn from the newest LOGITS row by the fused kernel (softmax, filter and the
        multinomial race in one pass; the probability row is never written)."""
        row = self._logits_buffer[:, self._current_seq_len - 1 - self._win_base:self._current_seq_len - self._win_base]
        B, V = row.shape[0], row.shape[2]
        noise = torch.empty((B, V), dtype=torch.float32, device=row.device).exponential_(1.0)
        res = ops.verify_fused(row[:, :0], row, torch.empty((B, 0), dtype=torch.int64, device=row.device),
                               torch.empty((B, 0), dtype=torch.float32, device=row.device), noise,
                               float(self._temperature), int(self._top_k or 0), float(self._top_p or 0.0))
        return res["token"].unsqueeze(1)

    def _generate_with_kvcache(self, prefix: torch.Tensor, gamma: int) -> torch.Tensor:
        """src/model_gpu.py:308-322: gamma x (forward -> sample).  While nobody has patched this module's
        ``sample`` the token comes straight from the logits row."""
        x = prefix if prefix.dtype == torch.long else prefix.to(torch.long)
        for _ in range(gamma):
            if globals()["sample"] is _ORIG_SAMPLE and x.is_cuda:
                self._forward_logits(x)
                nxt = self._sample_last_row()
            else:
                nxt = sample(self._forward_with_kvcache(x)).to(torch.long)
            x = torch.cat((x, nxt), dim=1)
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
        self._current_seq_len = max(min(end_pos, self.current_length), self._win_base)
        self._prob_done = min(self._prob_done, self._current_seq_len)

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
        have = 0 if self._logits_buffer is None else self._current_seq_len
        return {"current_length": self.current_length, "tracked_seq_len": self._current_seq_len,
                "prob_history_len": have, "logits_history_len": have, "max_length": self.max_length}

    def debug_row_sums(self, start: int, end: int) -> List[float]:
        if self._logits_buffer is None or self._current_seq_len <= 0:
            return []
        start = max(self._win_base, min(start, self._current_seq_len))
        end = max(start, min(end, self._current_seq_len))
        return self.prob_history[0, start:end, :].float().sum(dim=-1).tolist() if end > start else []

    def __len__(self) -> int:
        return self.current_length
