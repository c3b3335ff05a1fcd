"""InPlaceKVCache (duodecoding_b200.model_gpu): the in-place transformer cache behind KVCacheModel.rollback.

Runs on CPU (pure torch + transformers): logits through the in-place cache equal logits through the
reference's DynamicCache (src/model_gpu.py:362-383 crops that one), before and after a rollback, and a
rollback moves no cache bytes (same storage, shorter views)."""
import pytest
import torch

from duodecoding_b200.model_gpu import InPlaceKVCache, InPlaceLayer, RowWindow


def _tiny_llama():
    from transformers import LlamaConfig, LlamaForCausalLM
    torch.manual_seed(0)
    cfg = LlamaConfig(vocab_size=64, hidden_size=32, intermediate_size=64, num_hidden_layers=2,
                      num_attention_heads=4, num_key_value_heads=2, max_position_embeddings=256)
    return LlamaForCausalLM(cfg).eval()


@torch.no_grad()
def test_inplace_cache_matches_dynamic_cache_through_rollback():
    from transformers import DynamicCache
    model = _tiny_llama()
    ids = torch.randint(0, 64, (2, 9), generator=torch.Generator().manual_seed(1))
    ours, ref = InPlaceKVCache(), DynamicCache()
    a = model(ids[:, :5], past_key_values=ours, use_cache=True)
    b = model(ids[:, :5], past_key_values=ref, use_cache=True)
    torch.testing.assert_close(a.logits, b.logits)
    assert a.past_key_values is ours and ours.get_seq_length() == 5 and len(ours) == 2
    for t in range(5, 8):  # single-token decode steps
        a = model(ids[:, t:t + 1], past_key_values=ours, use_cache=True)
        b = model(ids[:, t:t + 1], past_key_values=ref, use_cache=True)
        torch.testing.assert_close(a.logits, b.logits)
    assert ours.get_seq_length() == 8
    ptrs = [layer._kbuf.data_ptr() for layer in ours.layers]
    ours.crop(6)
    ref.crop(6)
    assert ours.get_seq_length() == 6 and [k.shape[2] for k, _ in ours] == [6, 6]
    assert [layer._kbuf.data_ptr() for layer in ours.layers] == ptrs  # rollback: a length write, no bytes moved
    ours.crop(7)  # cropping past the length is a no-op (src/model_gpu.py:367-370 semantics)
    assert ours.get_seq_length() == 6
    a = model(ids[:, 6:9], past_key_values=ours, use_cache=True)  # re-forward a 3-token suffix after the rollback
    b = model(ids[:, 6:9], past_key_values=ref, use_cache=True)
    torch.testing.assert_close(a.logits, b.logits)
    fresh = model(ids, use_cache=False).logits[:, 6:9]
    torch.testing.assert_close(a.logits, fresh, rtol=1e-4, atol=1e-4)
    ours.crop(-2)
    assert ours.get_seq_length() == 7


def test_inplace_layer_grows_geometrically_and_keeps_contents():
    layer = InPlaceLayer()
    layer.initial_capacity = 4
    k = torch.arange(2 * 1 * 3 * 2, dtype=torch.float32).reshape(2, 1, 3, 2)
    layer.update(k, k + 100)
    assert layer._kbuf.shape[2] == 6
    k2 = torch.ones(2, 1, 5, 2)
    keys, values = layer.update(k2, k2)
    assert keys.shape[2] == 8 and layer._kbuf.shape[2] >= 8
    assert torch.equal(keys[:, :, :3], k) and torch.equal(values[:, :, :3], k + 100)
    assert torch.equal(keys[:, :, 3:], k2)


class _Owner:
    history_window = 4

    def __init__(self):
        self._logits_buffer = torch.arange(2 * 8 * 5, dtype=torch.float32).reshape(2, 8, 5)
        self._current_seq_len, self._win_base = 13, 7  # buffer rows 0..5 hold absolute rows 7..12

    def _window_tensor(self, kind):
        return self._logits_buffer[:, :self._current_seq_len - self._win_base]


def test_row_window_serves_absolute_rows_and_rejects_rows_that_left():
    o = _Owner()
    h = RowWindow(o, "logits")
    assert tuple(h.shape) == (2, 13, 5) and h.size(1) == 13 and h.dim() == 3
    assert torch.equal(h[:, 9:12, :3], o._logits_buffer[:, 2:5, :3])
    assert torch.equal(h[:, -1, :], o._logits_buffer[:, 5, :])
    assert torch.equal(h[:, 8, :2], o._logits_buffer[:, 1, :2])
    assert h[:, 5:5, :].shape[1] == 0  # empty slices never touch storage
    h[:, 10, :2] = -1.0  # callers write into the history (src/engine.py:948-954)
    assert float(o._logits_buffer[0, 3, 0]) == -1.0 and float(o._logits_buffer[1, 3, 1]) == -1.0
    with pytest.raises(IndexError):
        _ = h[:, 3:9, :]
    with pytest.raises(IndexError):
        _ = h[:, 6, :]
    with pytest.raises(IndexError):
        _ = h[:, 13, :]
