"""GPU tests of the reference-named mirror modules (duodecoding_b200.decoding_ops / utils / model_gpu /
communication): the reference's own test cases re-hosted on cuda, plus oracle comparisons."""
import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def mods():
    from duodecoding_b200 import communication, decoding_ops, model_gpu, utils
    return decoding_ops, utils, model_gpu, communication


class Cache:
    def __init__(self, probs, vocab_size):
        self.prob_history, self.vocab_size, self.device = probs, vocab_size, probs.device
        self.rollback_calls = []

    def rollback(self, end_pos):
        self.rollback_calls.append(end_pos)


def test_acceptance_matches_oracle(mods, dev):
    dops = mods[0]
    g = torch.Generator().manual_seed(3)
    B, L, V, P, gamma = 3, 12, 500, 6, 4
    q = torch.softmax(torch.randn(B, L, V, generator=g), -1)
    p = torch.softmax(torch.randn(B, L, V, generator=g), -1)
    x = torch.randint(0, V, (B, P + gamma), generator=g)
    r = torch.rand(B, gamma, generator=g) * 1.5
    vi, res = dops.verify_draft_sequence_result(Cache(q.to(dev), V), Cache(p.to(dev), V), x.to(dev), P, gamma,
                                                r=r.to(dev))
    want = oracle.compute_acceptance(q[:, P - 1:P - 1 + gamma], p[:, P - 1:P - 1 + gamma], x[:, P:P + gamma], P, r)
    assert isinstance(res.accepted_count, int) and isinstance(res.n, int)
    assert (res.accepted_count, res.n) == (want["accepted_count"], want["n"])
    assert torch.equal(res.accept_mask.cpu(), want["accept_mask"])
    torch.testing.assert_close(res.selected_target_p.cpu(), want["selected_target_p"], rtol=1e-6, atol=0)
    torch.testing.assert_close(res.selected_draft_p.cpu(), want["selected_draft_p"], rtol=1e-6, atol=0)


def test_reference_known_answers(mods, dev):  # test/test_temperature_sampling.py:79-208
    dops, utils = mods[0], mods[1]
    q = torch.tensor([[[0.0, 1.0, 0.0]]], device=dev)
    p = torch.tensor([[[0.5, 0.25, 0.25]]], device=dev)
    vi = dops.prepare_verification_inputs(Cache(q, 3), Cache(p, 3), torch.tensor([[7, 0]], device=dev), 1, 1)
    res = dops.compute_acceptance_result(vi, r=torch.tensor([[0.99]], device=dev))
    assert res.accepted_count == 1 and bool(res.accept_mask[0, 0])  # q = 0 < p: inf >= r accepts
    resid = dops.compute_residual_distribution(torch.tensor([[0.5, 0.3, 0.2]], device=dev),
                                               torch.tensor([[0.2, 0.5, 0.3]], device=dev))
    torch.testing.assert_close(resid.cpu(), torch.tensor([[1.0, 0.0, 0.0]]))
    same = torch.tensor([[0.25, 0.25, 0.5]], device=dev)
    assert dops.sample_reject_token(same, same).tolist() == [[0]]  # zero residual mass -> first argmax
    t = dops.sample_reject_token(torch.tensor([[0.1, 0.2, 0.7]], device=dev), torch.tensor([[0.3, 0.3, 0.4]], device=dev),
                                 output_device=torch.device("cpu"))
    assert t.device.type == "cpu" and t.tolist() == [[2]] and t.dtype == torch.int64
    hot = utils.norm_logits(torch.tensor([[[1.0, 3.0, 2.0]]], device=dev), 0.0, 0, 0.0)
    assert hot.dtype == torch.float32 and hot.tolist() == [[[0.0, 1.0, 0.0]]]
    filt = utils.top_k_top_p_filter(torch.tensor([[1.0, 4.0, 3.0, 2.0]], device=dev), top_k=2)
    assert torch.isinf(filt[0, [0, 3]]).all() and filt[0, 1] == 4.0 and filt[0, 2] == 3.0
    reb = utils.rebuild_topk_uniform_probs(torch.tensor([[0.5, 0.3, 0.1, 0.1]], device=dev), 2)
    torch.testing.assert_close(reb.cpu(), torch.tensor([[0.5, 0.3, 0.1, 0.1]]))


def test_sample_with_injected_noise_matches_oracle(mods, dev):
    utils = mods[1]
    g = torch.Generator().manual_seed(5)
    probs = torch.softmax(torch.randn(16, 3000, generator=g) * 2, -1)
    noise = torch.empty(16, 3000).exponential_(1.0, generator=g)
    assert torch.equal(utils.sample(probs.to(dev), noise=noise.to(dev)).cpu(), oracle.sample(probs, noise))
    drawn = utils.sample(probs.to(dev))  # device generator: a valid index of the support
    assert drawn.shape == (16, 1) and int(drawn.min()) >= 0 and int(drawn.max()) < 3000


def test_finalize_and_resolve_with_real_kernels(mods, dev):
    dops = mods[0]
    g = torch.Generator().manual_seed(9)
    V, P, gamma = 300, 4, 3
    q = torch.softmax(torch.randn(1, P + gamma, V, generator=g), -1).to(dev)
    x = torch.randint(0, V, (1, P + gamma), generator=g).to(dev)
    for n, accept in [(P + gamma - 1, True), (P, False)]:
        d, t = Cache(q, V), Cache(q.clone(), V)  # identical models
        out = dops.finalize_verification(d, t, x, P, gamma, n)
        assert out.shape == (1, n + 2) and torch.equal(out[:, :n + 1], x[:, :n + 1])
        assert d.rollback_calls == [n + 1] and t.rollback_calls == [n + 2 if accept else n + 1]
        assert 0 <= int(out[0, -1]) < V
    d, t = Cache(q, V), Cache(q.clone(), V)
    acc, n, tok, all_acc = dops.resolve_stage_verification(d, t, x, P, gamma, output_device=dev)
    assert (acc, n, all_acc) == (gamma, P + gamma - 1, True)  # p == q: every draft is accepted
    assert d.rollback_calls == [P + gamma] and t.rollback_calls == [P + gamma + 1]


def test_communication_helpers_match_oracle(mods, dev):
    comm = mods[3]
    g = torch.Generator().manual_seed(11)
    p1 = torch.softmax(torch.randn(700, generator=g), -1)
    torch.testing.assert_close(comm.CommunicationSimulator._apply_top_k_compression(p1.to(dev), 20).cpu(),
                               oracle.apply_top_k_compression(p1, 20))
    p3 = torch.softmax(torch.randn(2, 3, 700, generator=g), -1)
    torch.testing.assert_close(comm.compress_rebuild_probs(p3.to(dev), 50).cpu(), oracle.compress_rebuild_probs(p3, 50),
                               rtol=1e-5, atol=1e-9)
    comp = oracle.apply_top_k_compression(p1, 20).reshape(1, -1)
    torch.testing.assert_close(comm.rebuild_full_probs(comp.to(dev)).cpu(), oracle.rebuild_full_probs(comp),
                               rtol=1e-5, atol=1e-9)


def _tiny_llama(dev):
    from transformers import LlamaConfig, LlamaForCausalLM
    torch.manual_seed(0)
    cfg = LlamaConfig(vocab_size=64, hidden_size=32, intermediate_size=64, num_hidden_layers=2,
                      num_attention_heads=4, num_key_value_heads=4, max_position_embeddings=128)
    return cfg, LlamaForCausalLM(cfg).to(dev).eval()


def test_kvcache_model_rollback(mods, dev):  # test/test_model_gpu_rollback.py
    mg = mods[2]
    cfg, model = _tiny_llama(dev)
    cache = mg.KVCacheModel(model, temperature=0, top_k=0, top_p=0)
    cache.vocab_size = cfg.vocab_size
    prefix = torch.tensor([[1, 2, 3, 4, 5]], device=dev)
    x = cache.generate(prefix, 2)
    assert tuple(x.shape) == (1, 7) and cache.current_length == 6 and cache._current_seq_len == 6
    cache.rollback(7)  # must not extend past the real cache length
    assert cache.current_length == 6 and cache._current_seq_len == 6 and cache.prob_history.shape[1] == 6
    assert cache.seq_lens.tolist() == [6] and len(cache) == 6
    cache.rollback(4)
    assert cache.current_length == 4 and cache.prob_history.shape[1] == 4 and cache.seq_lens.tolist() == [4]
    # contents after rollback + re-forward == a fresh forward of the truncated prefix (unpinned upstream)
    again = cache._forward_with_kvcache(x[:, :6])
    fresh = mg.KVCacheModel(model, temperature=0, top_k=0, top_p=0)
    want = fresh._forward_with_kvcache(x[:, :6])
    torch.testing.assert_close(again, want)
    torch.testing.assert_close(cache.logits_history, fresh.logits_history, rtol=1e-4, atol=1e-4)
    with pytest.raises(ValueError):
        mg.KVCacheModel(model)._forward_with_kvcache(torch.tensor([[1, 99]], device=dev))
    with pytest.raises(ValueError):
        _ = mg.KVCacheModel(model).prob_history


def test_kvcache_history_rows_match_oracle_and_rebuilt_generation(mods, dev):
    mg = mods[2]
    cfg, model = _tiny_llama(dev)
    cache = mg.KVCacheModel(model, temperature=0.8, top_k=0, top_p=0.9)
    x, rebuilt = cache.generate_with_rebuilt_topk(torch.tensor([[3, 1, 4, 1, 5]], device=dev), 3, 8)
    assert tuple(x.shape) == (1, 8) and tuple(rebuilt.shape) == (1, 3, 64)
    want = oracle.norm_logits(cache.logits_history.float().cpu(), 0.8, 0, 0.9)
    torch.testing.assert_close(cache.prob_history.float().cpu(), want, rtol=1e-4, atol=1e-7)
    torch.testing.assert_close(rebuilt.sum(-1).cpu(), torch.ones(1, 3), rtol=1e-5, atol=1e-6)
    cache.prob_history[:, 2, :10] = 0.0  # callers write into the history (src/engine.py:948-954)
    assert float(cache._prob_buffer[0, 2, :10].abs().sum()) == 0.0


def test_static_kv_cache_crop_is_a_length_write(mods, dev):
    mg = mods[2]
    kv = mg.StaticKVCache(num_layers=2, batch=3, heads=2, max_len=16, head_dim=4, dtype=torch.float32, device=dev)
    for layer in range(2):
        kv.append(layer, torch.ones(3, 2, 5, 4, device=dev), torch.ones(3, 2, 5, 4, device=dev))
    ptr = kv.keys[0].data_ptr()
    assert kv.get_seq_length() == 5 and kv.seq_lens.tolist() == [5, 5, 5]
    kv.crop(3)
    assert kv.get_seq_length() == 3 and kv.seq_lens.tolist() == [3, 3, 3] and kv.keys[0].data_ptr() == ptr
    assert [k.shape[2] for k, _ in kv] == [3, 3]
    kv.crop(torch.tensor([1, 9, 2], device=dev))  # ragged per-sequence lengths
    assert kv.seq_lens.tolist() == [1, 3, 2]
