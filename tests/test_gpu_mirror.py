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
    kv = cache._past_key_values
    assert isinstance(kv, mg.InPlaceKVCache) and kv.seq_lens.tolist() == [6] and len(cache) == 6
    ptrs = [layer._kbuf.data_ptr() for layer in kv.layers]
    cache.rollback(4)
    assert cache.current_length == 4 and cache.prob_history.shape[1] == 4 and kv.seq_lens.tolist() == [4]
    assert [layer._kbuf.data_ptr() for layer in kv.layers] == ptrs  # a length write: no K/V byte moved
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


def test_inplace_kv_cache_ragged_crop_clamps_device_lengths(mods, dev):
    mg = mods[2]
    kv = mg.InPlaceKVCache()
    for layer in range(2):
        kv.update(torch.ones(3, 2, 5, 4, device=dev), torch.ones(3, 2, 5, 4, device=dev), layer)
    ptr = kv.layers[0]._kbuf.data_ptr()
    assert kv.get_seq_length() == 5 and kv.seq_lens.tolist() == [5, 5, 5]
    kv.crop(3)
    assert kv.get_seq_length() == 3 and kv.seq_lens.tolist() == [3, 3, 3] and kv.layers[0]._kbuf.data_ptr() == ptr
    assert [k.shape[2] for k, _ in kv] == [3, 3]
    kv.crop(torch.tensor([1, 9, 2], device=dev))  # ragged per-sequence lengths (cache_seqlens of attention kernels)
    assert kv.seq_lens.tolist() == [1, 3, 2]


# ------------------------------------------------------------------ lazy histories, window, fused routes
class _NoModel:
    """What KVCacheModel's constructor needs from a model, for tests that load logit rows by hand."""

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
    """Stand-in transformer cache: a length with crop / get_seq_length (src/model_gpu.py:23-26)."""

    def __init__(self, n):
        self.n, self.crops = n, []

    def crop(self, end):
        self.crops.append(end)
        self.n = min(self.n, end)

    def get_seq_length(self, layer_idx=0):
        return self.n


def _loaded_cache(mg, logits, T, k, p, **kw):
    """A KVCacheModel whose history holds the given [B, L, V] logit rows."""
    cache = mg.KVCacheModel(_NoModel(logits.shape[-1], logits.device), temperature=T, top_k=k, top_p=p, **kw)
    cache._store_rows(logits, 0)
    cache._past_key_values = _Len(logits.shape[1])
    return cache


def test_probabilities_are_lazy_and_one_launch(mods, dev):
    mg = mods[2]
    g = torch.Generator().manual_seed(5)
    x = (torch.randn(3, 9, 700, generator=g) * 3).to(dev)
    cache = _loaded_cache(mg, x, 0.8, 0, 0.9)
    assert cache._prob_buffer is None  # nothing V-sized was normalised by the store
    assert cache.logits_rows(4, 9).data_ptr() == cache._logits_buffer[:, 4:9].data_ptr()
    want = oracle.norm_logits(x.cpu(), 0.8, 0, 0.9)
    torch.testing.assert_close(cache.prob_history.cpu(), want, rtol=1e-4, atol=1e-7)
    cache.prob_history[:, 2, :10] = 0.0  # callers write into the history (src/engine.py:948-954)
    assert float(cache._prob_buffer[0, 2, :10].abs().sum()) == 0.0
    cache._store_rows(x[:, :2], 9)  # two more rows: only they are normalised on the next read
    assert float(cache.prob_history[0, 2, :10].abs().sum()) == 0.0 and cache.prob_history.shape[1] == 11
    cache.rollback(7)
    assert cache.prob_history.shape[1] == 7 and cache._past_key_values.crops == [7]


def test_history_window_keeps_memory_flat_and_serves_recent_rows(mods, dev):
    mg = mods[2]
    g = torch.Generator().manual_seed(6)
    V, W = 512, 4
    x = (torch.randn(2, 23, V, generator=g) * 2).to(dev)
    cache = _loaded_cache(mg, x[:, :5], 1.0, 0, 0.0, history_window=W)
    for t in range(5, 23):
        cache._store_rows(x[:, t:t + 1], t)
        cache._past_key_values.n = t + 1
    assert cache._logits_buffer.shape[1] == 2 * W and cache._current_seq_len == 23
    hist = cache.prob_history
    assert tuple(hist.shape) == (2, 23, V)
    want = oracle.norm_logits(x[:, 20:23].cpu(), 1.0, 0, 0.0)
    torch.testing.assert_close(hist[:, 20:23, :].cpu(), want, rtol=1e-4, atol=1e-7)
    torch.testing.assert_close(hist[:, -1, :V].cpu(), want[:, -1], rtol=1e-4, atol=1e-7)
    torch.testing.assert_close(cache.logits_history[:, 21, :].cpu(), x[:, 21].cpu())
    hist[:, 22, :7] = 0.0
    assert float(cache.prob_history[:, 22, :7].abs().sum()) == 0.0
    with pytest.raises(IndexError):
        _ = hist[:, 3, :]
    # B=256, V=151936 (config 5) stays under 2 GB with a window: 256 * 2W * V * 2 bytes
    big = mg.KVCacheModel(_NoModel(151936, dev), history_window=8)
    big._ensure_buffer_size(256, 1, dev, torch.bfloat16)
    assert big._logits_buffer.numel() * 2 < 2 * 1024 ** 3 and big._prob_buffer is None


def _fused_case(seed, B, P, gamma, V, sigma, dtype=torch.float32):
    g = torch.Generator().manual_seed(seed)
    tl = torch.randn(B, P + gamma, V, generator=g) * 3
    dl = tl[:, :P + gamma - 1] + sigma * torch.randn(B, P + gamma - 1, V, generator=g)
    tl, dl = tl.to(dtype), dl.to(dtype)
    q = oracle.norm_logits(dl[:, P - 1:].float(), 1.0, 0, 0.9)
    toks = torch.stack([oracle.sample(q[:, j], torch.empty(B, V).exponential_(1, generator=g))[:, 0]
                        for j in range(gamma)], 1)
    x = torch.cat((torch.randint(0, V, (B, P), generator=g), toks), 1)
    r = torch.rand(B, gamma, generator=g)
    noise = torch.empty(B, V).exponential_(1, generator=g)
    return dl, tl, x, r, noise


def _row0_expectation(dl, tl, x, r, noise, P, gamma):
    """The reference's flow on the CPU oracle: sequence 0's accepted count positions every sequence."""
    q = oracle.norm_logits(dl[:, P - 1:P - 1 + gamma].float(), 1.0, 0, 0.9)
    p = oracle.norm_logits(tl[:, P - 1:P + gamma].float(), 1.0, 0, 0.9)
    acc = oracle.compute_acceptance(q, p[:, :gamma], x[:, P:P + gamma], P, r)
    a = acc["accepted_count"]
    tok = oracle.sample_accept_token(p[:, gamma], noise) if a == gamma else \
        oracle.sample_reject_token(p[:, a], q[:, a], noise)
    return a, P + a - 1, tok


@pytest.mark.parametrize("B,V,dtype", [(1, 4096, torch.float32), (5, 32000, torch.float32), (3, 8192, torch.bfloat16)])
def test_fused_routes_match_the_reference_flow(mods, dev, B, V, dtype):
    """resolve_stage_verification and verify_draft_sequence -> finalize_verification on KVCacheModel caches run
    one fused kernel on the logit rows; results equal the reference flow (batch rule included) on the oracle."""
    from unittest.mock import patch
    dops, mg = mods[0], mods[2]
    P, gamma = 6, 4
    for seed, sigma in [(41, 0.4), (42, 1.5), (43, 0.0)]:
        dl, tl, x, r, noise = _fused_case(seed, B, P, gamma, V, sigma, dtype)
        a, n, tok = _row0_expectation(dl, tl, x, r, noise, P, gamma)
        for route in ("resolve", "engine"):
            dc = _loaded_cache(mg, dl.to(dev), 1.0, 0, 0.9)
            tc = _loaded_cache(mg, tl.to(dev), 1.0, 0, 0.9)
            with patch.object(torch, "rand", lambda *s, **kw: r.to(dev)), \
                    patch.object(torch.Tensor, "exponential_", lambda self, *a_, **kw: self.copy_(noise.to(dev))):
                if route == "resolve":
                    got = dops.resolve_stage_verification(dc, tc, x.to(dev), P, gamma, output_device=dev)
                    ga, gn, gt, gall = got
                else:
                    ga, gn = dops.verify_draft_sequence(dc, tc, x.to(dev), P, gamma, transfer_mode="none",
                                                        send_reject_message=False)
                    assert tc._fused_token is not None
                    out = dops.finalize_verification(dc, tc, x.to(dev), P, gamma, gn)
                    assert tuple(out.shape) == (B, gn + 2) and torch.equal(out[:, :gn + 1].cpu(), x[:, :gn + 1])
                    gt, gall = out[:, -1:], gn >= P + gamma - 1
            assert dc._prob_buffer is None and tc._prob_buffer is None  # no probability row was materialised
            assert isinstance(ga, int) and isinstance(gn, int) and (ga, gn) == (a, n) and gall == (a == gamma)
            assert torch.equal(gt.cpu().reshape(-1), tok.reshape(-1)), (route, seed)
            assert dc._past_key_values.crops == [n + 1] and tc._past_key_values.crops == [n + 2 if a == gamma else n + 1]


def test_fused_route_steps_aside_for_overrides_and_patched_samplers(mods, dev):
    from unittest.mock import patch
    dops, mg = mods[0], mods[2]
    P, gamma, V = 5, 3, 4096
    dl, tl, x, r, noise = _fused_case(44, 1, P, gamma, V, 0.8)
    dc, tc = _loaded_cache(mg, dl.to(dev), 1.0, 0, 0.9), _loaded_cache(mg, tl.to(dev), 1.0, 0, 0.9)
    with patch("duodecoding_b200.decoding_ops.sample", lambda probs, *a, **k: torch.full((probs.shape[0], 1), 7, device=dev)):
        _, _, t, _ = dops.resolve_stage_verification(dc, tc, x.to(dev), P, gamma, output_device=dev)
    assert t.tolist() == [[7]] and tc._prob_buffer is not None  # the mocked sampler saw probability rows
    dc, tc = _loaded_cache(mg, dl.to(dev), 1.0, 0, 0.9), _loaded_cache(mg, tl.to(dev), 1.0, 0, 0.9)
    override = dc.prob_history.clone()
    dops.resolve_stage_verification(dc, tc, x.to(dev), P, gamma, output_device=dev, draft_probs_override=override)
    assert tc._prob_buffer is not None


def test_golden_resolve_and_finalize_replayed_on_the_gpu(golden, mods, dev):
    """The reference-generated resolve/* and finalize/* fixtures through the GPU mirrors (probability-space
    operators on plain history tensors), bit-exact."""
    from unittest.mock import patch
    import duodecoding_b200.ops as ops
    dops = mods[0]
    Pm, Qm = torch.from_numpy(golden["finalize/P"]).to(dev), torch.from_numpy(golden["finalize/Q"]).to(dev)
    x = torch.from_numpy(golden["finalize/x"]).to(dev)
    noise = torch.from_numpy(golden["finalize/noise"]).to(dev)
    V = Pm.shape[-1]
    with patch.object(ops, "draw_noise", lambda rows, v, d, generator=None: noise[:rows, :v]):
        for name in ("reject", "accept"):
            P_, g_, n_ = (int(v) for v in golden[f"finalize/{name}/args"])
            a, t = Cache(Qm.clone(), V), Cache(Pm.clone(), V)
            y = dops.finalize_verification(a, t, x, P_, g_, n_)
            assert torch.equal(y.cpu(), torch.from_numpy(golden[f"finalize/{name}/out"]))
            assert a.rollback_calls + t.rollback_calls == golden[f"finalize/{name}/rollbacks"].tolist()
        for name in ("mixed", "all"):
            rr = torch.from_numpy(golden[f"resolve/{name}/r"]).to(dev)
            a, t = Cache(Qm.clone(), V), Cache(Pm.clone(), V)
            with patch.object(torch, "rand", lambda *s, **kw: rr.clone()):
                ac, n, tk, alla = dops.resolve_stage_verification(a, t, x, 8, 4, output_device=dev)
            assert [ac, n, int(tk), int(alla)] == golden[f"resolve/{name}/result"].tolist()
            assert a.rollback_calls + t.rollback_calls == golden[f"resolve/{name}/rollbacks"].tolist()


class _Comm:
    def __init__(self):
        self.calls = []

    def transfer(self, *args):
        self.calls.append(("transfer",) + tuple(a.shape if isinstance(a, torch.Tensor) else a for a in args))

    def send_reject_message(self, link):
        self.calls.append(("reject", link))


def test_verify_draft_sequence_comm_accounting_and_empty_history(mods, dev):  # src/decoding_ops.py:223-328
    from unittest.mock import patch
    dops = mods[0]
    g = torch.Generator().manual_seed(9)
    B, L, V, P, gamma = 1, 10, 300, 5, 4
    q = torch.softmax(torch.randn(B, L, V, generator=g), -1).to(dev)
    p = torch.softmax(torch.randn(B, L, V, generator=g), -1).to(dev)
    x = torch.randint(0, V, (B, P + gamma), generator=g).to(dev)
    r = torch.rand(B, gamma, generator=g)
    want = oracle.compute_acceptance(q[:, P - 1:P - 1 + gamma].cpu(), p[:, P - 1:P - 1 + gamma].cpu(),
                                     x[:, P:P + gamma].cpu(), P, r)
    for mode in ("none", "serial", "batch_before"):
        comm, metrics = _Comm(), {"draft_generated_tokens": 0, "draft_accepted_tokens": 0}
        with patch.object(torch, "rand", lambda *s, **kw: r.to(dev)):
            ac, n = dops.verify_draft_sequence(Cache(q, V), Cache(p, V), x, P, gamma, comm_simulator=comm,
                                               comm_link="edge_cloud", transfer_mode=mode, decoding_metrics=metrics)
        assert (ac, n) == (want["accepted_count"], want["n"])
        assert metrics == {"draft_generated_tokens": gamma, "draft_accepted_tokens": n - P + 1}
        rejected = ac < gamma
        kinds = [c[0] for c in comm.calls]
        assert kinds.count("reject") == (1 if rejected else 0)
        assert kinds.count("transfer") == {"none": 0, "batch_before": 1, "serial": ac + (1 if rejected else 0)}[mode]
    # nothing to verify: early return, no traffic, no metrics (reference: `if actual_gamma <= 0: return 0, prefix_len - 1`)
    comm, metrics = _Comm(), {"draft_generated_tokens": 0, "draft_accepted_tokens": 0}
    ac, n = dops.verify_draft_sequence(Cache(q[:, :4], V), Cache(p, V), x, P, gamma, comm_simulator=comm,
                                       transfer_mode="batch_before", decoding_metrics=metrics)
    assert (ac, n) == (0, P - 1) and comm.calls == [] and metrics["draft_generated_tokens"] == 0


def test_cuhlm_helpers_match_reference_goldens_and_oracle(mods, dev):
    """CUHLM.calculate_uncertainty / _calculate_compressed_vocab_size (src/communication.py:541-575, :681-755)."""
    import os
    from unittest.mock import patch

    import numpy as np

    import duodecoding_b200.ops as ops
    comm = mods[3]
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "cuhlm_v1.npz"))
    for ci in range(3):  # the reference's own outputs, with its temperatures and multinomial noise injected
        pre = f"unc/{ci}/"
        M, theta_max, draft = z[pre + "params"]
        u, noise = torch.from_numpy(z[pre + "u"]).to(dev), torch.from_numpy(z[pre + "noise"]).to(dev)
        with patch.object(torch, "rand", lambda *a, **k: u.clone()), \
                patch.object(ops, "draw_noise", lambda rows, v, d, generator=None: noise):
            got = comm.CUHLM.calculate_uncertainty(torch.from_numpy(z[pre + "logits"]).to(dev), M=int(M),
                                                   theta_max=float(theta_max), draft_token=int(draft))
        assert got == float(z[pre + "value"][0]), ci
    for ci in range(5):
        pre = f"kstar/{ci}/"
        unc, theta, draft = z[pre + "params"]
        probs = torch.from_numpy(z[pre + "probs"]).to(dev)
        sim = comm.CUHLM(vocab_size=len(probs))
        got = sim._calculate_compressed_vocab_size(float(unc), probs, theta=float(theta),
                                                   draft_token=None if draft < 0 else int(draft))
        assert got == int(z[pre + "k"][0]), (ci, got)
    # a full-size vocabulary against the oracle (the reference's O(V^2) loop would take hours here)
    g = torch.Generator().manual_seed(8)
    V = 32000
    logits = torch.randn(V, generator=g) * 3
    temps = torch.clamp(torch.rand(20, generator=g) * 2.0, min=1e-6)
    noise = torch.empty(20, V).exponential_(1, generator=g)
    want = oracle.calculate_uncertainty_tokens(logits, temps, noise)
    got = comm.uncertainty_tokens(logits.to(dev), temps.to(dev), noise.to(dev))
    assert torch.equal(got.cpu(), want.reshape(-1))
    probs = torch.softmax(logits, -1)
    for unc, theta in ((0.3, 0.1), (0.9, 0.02)):
        want_k = oracle.compressed_vocab_size(unc, probs, V, theta, 5)
        assert comm.compressed_vocab_size(unc, probs.to(dev), V, theta, 5) == want_k
    assert comm.CUHLM.calculate_uncertainty(None) == 1.0
    assert comm.CUHLM(vocab_size=10)._calculate_compressed_vocab_size(0.5, probs[:7].to(dev)) == 1  # length mismatch rule


def test_rank_helpers_match_reference_goldens_and_oracle(dev):
    """duodecoding_b200.baselines (one kernel + one read-back for all accepted positions) against the reference
    outputs in tests/golden/ranks_v1.npz and, on rows full of equal values, against the oracle's canonical order."""
    import os

    import numpy as np

    import oracle
    from duodecoding_b200 import baselines as B

    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "ranks_v1.npz"))
    for ci in sorted({k.split("/")[1] for k in z.files}):
        pre = f"ranks/{ci}/"
        prefix, accepted, k = (int(v) for v in z[pre + "params"])
        probs, x = torch.from_numpy(z[pre + "probs"]).to(dev), torch.from_numpy(z[pre + "x"]).to(dev)
        vr, it, tr = [7], [True], [7]  # the histories are appended to
        B._record_accepted_token_ranks(stage_probs=probs, x=x, prefix_len=prefix, accepted_count=accepted,
                                       transfer_top_k=k, vocab_rank_history=vr, in_transfer_topk_history=it,
                                       transfer_topk_rank_history=tr)
        assert vr[1:] == z[pre + "vocab_rank"].tolist() and it[1:] == z[pre + "in_topk"].tolist(), ci
        assert tr[1:] == z[pre + "topk_rank"].tolist(), ci
        if accepted:
            tok = int(x[0, prefix])
            rank = B._compute_token_vocab_rank(probs[0, prefix - 1], tok)
            assert rank == vr[1]
            assert B._compute_transfer_topk_rank(probs[0, prefix - 1], tok, k, rank) == (it[1], tr[1])
    # equal values (bf16 grid), bf16 storage, a strided history, k at the tie
    g = torch.Generator().manual_seed(5)
    pf = torch.softmax((torch.randn(1, 9, 4096, generator=g) * 2).bfloat16().float(), -1)
    for probs in (pf, pf.bfloat16()):
        x = torch.randint(0, 4096, (1, 10), generator=g)
        x[0, 3] = int(torch.argsort(pf[0, 2], descending=True, stable=True)[40])
        for k in (None, 41, 40, 300):
            want = ([], [], [])
            oracle.record_accepted_token_ranks(stage_probs=probs.float(), x=x, prefix_len=2, accepted_count=6,
                                               transfer_top_k=k, vocab_rank_history=want[0],
                                               in_transfer_topk_history=want[1], transfer_topk_rank_history=want[2])
            got = ([], [], [])
            B._record_accepted_token_ranks(stage_probs=probs.to(dev), x=x.to(dev), prefix_len=2, accepted_count=6,
                                           transfer_top_k=k, vocab_rank_history=got[0],
                                           in_transfer_topk_history=got[1], transfer_topk_rank_history=got[2])
            assert got == want, (probs.dtype, k)
    none = ([], [], [])
    B._record_accepted_token_ranks(stage_probs=None, x=x.to(dev), prefix_len=2, accepted_count=3, transfer_top_k=8,
                                   vocab_rank_history=none[0], in_transfer_topk_history=none[1],
                                   transfer_topk_rank_history=none[2])
    assert none == ([], [], [])
