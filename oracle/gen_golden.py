#!/usr/bin/env python
"""Generate tests/golden/*.npz by running the UNMODIFIED reference.

Run in the build container only (needs /root/reference):

    python oracle/gen_golden.py

The reference modules are imported from /root/reference as they lie; the only
interposition is ``torch.multinomial`` -> ``argmax(p / noise)`` with a
pre-drawn Exp(1) tensor, which is what ATen's multinomial does internally
(checked below against the real ``torch.multinomial`` with a shared generator
state).  Inputs are seeded fp32 tensors without duplicate values in a row, so
the reference's unstable sort/top-k cannot make the outputs ambiguous.
Nothing here is used at run time on the GPU box; the fixtures are.
"""

from __future__ import annotations

import contextlib
import os
import sys
from types import SimpleNamespace

import numpy as np
import torch

REF = os.environ.get("DUODEC_REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _import_reference():
    sys.path.insert(0, REF)
    import src.utils as U  # noqa: E402
    import src.decoding_ops as D  # noqa: E402
    import src.decoding_types as T  # noqa: E402
    import src.communication as C  # noqa: E402
    import src.proposal_utils as P  # noqa: E402
    import src.model_gpu as M  # noqa: E402

    return SimpleNamespace(U=U, D=D, T=T, C=C, P=P, M=M)


@contextlib.contextmanager
def injected_noise(noise_rows):
    """Route torch.multinomial(p, 1) to argmax(p / noise) for queued noise rows."""
    real = torch.multinomial
    queue = list(noise_rows)

    def fake(p, num_samples=1, replacement=False, *, generator=None, out=None):
        assert num_samples == 1
        e = queue.pop(0)
        return torch.argmax(p / e.to(p.dtype), dim=-1, keepdim=True)

    torch.multinomial = fake
    try:
        yield
    finally:
        torch.multinomial = real


def check_multinomial_is_race():
    for s in range(8):
        p = torch.softmax(torch.randn(3, 777, generator=torch.Generator().manual_seed(100 + s)) * 3, -1)
        a = torch.multinomial(p, 1, generator=torch.Generator().manual_seed(s))
        e = torch.empty_like(p).exponential_(1, generator=torch.Generator().manual_seed(s))
        assert torch.equal(a, torch.argmax(p / e, -1, keepdim=True)), "multinomial != race"


def no_dup_rows(x: torch.Tensor) -> bool:
    flat = x.reshape(-1, x.shape[-1])
    return all(torch.unique(row).numel() == row.numel() for row in flat)


def gen_logits(gen, shape, scale=3.0):
    while True:
        x = torch.randn(shape, generator=gen) * scale
        if no_dup_rows(x):
            return x


class _FakeCache:
    """Prob-history holder with the rollback recorder the reference tests use."""

    def __init__(self, probs, vocab_size, device=torch.device("cpu")):
        self._p = probs
        self.vocab_size = vocab_size
        self.device = device
        self.rollback_calls = []

    @property
    def prob_history(self):
        return self._p

    def rollback(self, end_pos):
        self.rollback_calls.append(int(end_pos))


def main():
    os.makedirs(OUT, exist_ok=True)
    check_multinomial_is_race()
    R = _import_reference()
    g = torch.Generator().manual_seed(20261104)
    out = {}

    # ---- norm_logits over the warp grid -------------------------------------------------
    warp_grid = [
        (1.0, 0, 0.0),
        (0.7, 0, 0.0),
        (1.0, 0, 0.9),
        (1.3, 0, 0.5),
        (1.0, 50, 0.0),
        (0.8, 40, 0.95),
        (0.0, 0, 0.0),
        (1.0, 5000, 0.9),  # top_k > V is clamped
    ]
    shapes = [(2, 3, 257), (1, 5, 1000), (3, 1, 4096)]
    for si, shp in enumerate(shapes):
        x = gen_logits(g, shp)
        out[f"norm/{si}/logits"] = x.numpy()
        for wi, (T, k, p) in enumerate(warp_grid):
            y = R.U.norm_logits(x, T, k, p)
            out[f"norm/{si}/{wi}/params"] = np.array([T, k, p], dtype=np.float64)
            out[f"norm/{si}/{wi}/probs"] = y.numpy()
    # 2-D input (the reference accepts [B,V] too)
    x = gen_logits(g, (4, 513))
    out["norm2d/logits"] = x.numpy()
    out["norm2d/probs"] = R.U.norm_logits(x, 0.9, 20, 0.8).numpy()

    # ---- sample / max_fn ----------------------------------------------------------------
    p = torch.softmax(gen_logits(g, (6, 777)), -1)
    p[1] = 0.0  # zero-mass row -> argmax fallback (index 0)
    p[2, 5] = float("nan")
    p[3, 7] = float("inf")
    p[4] = -p[4]  # all negative -> clamp -> zero mass
    e = torch.empty(6, 777).exponential_(1, generator=g)
    with injected_noise([e]):
        idx = R.U.sample(p)
    out["sample/probs"] = p.numpy()
    out["sample/noise"] = e.numpy()
    out["sample/idx"] = idx.numpy()

    d = gen_logits(g, (5, 300), scale=0.01)
    d[1] = -d[1].abs()  # no positive entry -> one-hot at argmax
    d[2] = 0.0
    d[3, 4] = float("nan")
    d[3, 9] = float("inf")
    out["maxfn/x"] = d.numpy()
    out["maxfn/y"] = R.U.max_fn(d).numpy()

    # ---- rebuild_topk_probs -------------------------------------------------------------
    pr = torch.softmax(gen_logits(g, (4, 1000)), -1)
    out["rebuild/probs"] = pr.numpy()
    for k in (1, 7, 64, 999, 1000, 0):
        out[f"rebuild/k{k}"] = R.U.rebuild_topk_uniform_probs(pr, k).numpy()
    # nucleus-filtered input: fewer than k non-zero entries
    prn = R.U.norm_logits(gen_logits(g, (3, 1, 1000)), 1.0, 0, 0.6)[:, 0]
    out["rebuild_sparse/probs"] = prn.numpy()
    out["rebuild_sparse/k512"] = R.U.rebuild_topk_uniform_probs(prn, 512).numpy()

    # ---- communication helpers ----------------------------------------------------------
    CS = R.C.CommunicationSimulator
    v1 = torch.softmax(gen_logits(g, (1, 600)), -1)[0]
    out["comm/probs1d"] = v1.numpy()
    for k in (1, 32, 600):
        out[f"comm/topk{k}"] = CS._apply_top_k_compression(v1, k).numpy()
    comp = CS._apply_top_k_compression(v1, 32)
    out["comm/rebuild_full32"] = CS.rebuild_full_probs(comp).numpy()
    p3 = torch.softmax(gen_logits(g, (2, 3, 400)), -1)
    out["comm/probs3d"] = p3.numpy()
    out["comm/compress_rebuild16"] = CS.compress_rebuild_probs(p3, 16).numpy()
    out["comm/rebuild_full3d"] = CS.rebuild_full_probs(
        torch.where(p3 >= torch.topk(p3, 16, dim=-1)[0][..., -1:], p3, torch.zeros_like(p3))
    ).numpy()

    # ---- acceptance + token sampling on probability histories ---------------------------
    cases = []
    for ci, (B, gam, V, T, k, tp, sig) in enumerate(
        [
            (1, 4, 500, 1.0, 0, 0.0, 0.5),
            (1, 4, 500, 1.0, 0, 0.9, 1.0),
            (3, 5, 1024, 0.8, 30, 0.0, 1.0),
            (4, 4, 2048, 1.0, 0, 0.9, 0.3),
            (2, 6, 777, 0.0, 0, 0.0, 0.5),
            (4, 3, 333, 1.2, 10, 0.7, 2.0),
            (8, 4, 4096, 1.0, 0, 0.9, 0.5),
        ]
    ):
        tl = gen_logits(g, (B, gam + 1, V))
        dl = tl[:, :gam] + sig * torch.randn(B, gam, V, generator=g)
        P = R.U.norm_logits(tl, T, k, tp)
        Q = R.U.norm_logits(dl, T, k, tp)
        # draft tokens ~ q via the race sampler
        toks = torch.zeros(B, gam, dtype=torch.long)
        for j in range(gam):
            ej = torch.empty(B, V).exponential_(1, generator=g)
            with injected_noise([ej]):
                toks[:, j] = R.U.sample(Q[:, j])[:, 0]
        r = torch.rand(B, gam, generator=g)
        noise = torch.empty(B, V).exponential_(1, generator=g)
        acc = np.zeros(B, dtype=np.int64)
        tok = np.zeros(B, dtype=np.int64)
        mask = np.zeros((B, gam), dtype=bool)
        selp = np.zeros((B, gam), dtype=np.float32)
        selq = np.zeros((B, gam), dtype=np.float32)
        nvals = np.zeros(B, dtype=np.int64)
        prefix_len = 7
        for b in range(B):
            vi = R.T.VerificationInputs(
                draft_probs_batch=Q[b : b + 1],
                target_probs_batch=P[b : b + 1, :gam],
                draft_tokens=toks[b : b + 1],
                draft_token_indices=toks[b : b + 1].unsqueeze(-1),
                prefix_len=prefix_len,
                gamma=gam,
                actual_gamma=gam,
                max_idx=prefix_len + gam - 1,
            )
            res = R.D.compute_acceptance_result(vi, r=r[b : b + 1])
            acc[b] = res.accepted_count
            nvals[b] = res.n
            mask[b] = res.accept_mask[0].numpy()
            selp[b] = res.selected_target_p[0].float().numpy()
            selq[b] = res.selected_draft_p[0].float().numpy()
            plan = R.D.build_rollback_plan(prefix_len, gam, res.n)
            with injected_noise([noise[b : b + 1]]):
                if plan.all_accepted:
                    t = R.D.sample_accept_token(P[b : b + 1, -1, :V])
                else:
                    off = res.n - (prefix_len - 1)
                    t = R.D.sample_reject_token(P[b : b + 1, off, :V], Q[b : b + 1, off, :V])
            tok[b] = int(t)
        pre = f"verify/{ci}/"
        out[pre + "params"] = np.array([T, k, tp, prefix_len], dtype=np.float64)
        out[pre + "target_logits"] = tl.numpy()
        out[pre + "draft_logits"] = dl.numpy()
        out[pre + "tokens"] = toks.numpy()
        out[pre + "r"] = r.numpy()
        out[pre + "noise"] = noise.numpy()
        out[pre + "accepted_len"] = acc
        out[pre + "n"] = nvals
        out[pre + "token"] = tok
        out[pre + "accept_mask"] = mask
        out[pre + "selected_target_p"] = selp
        out[pre + "selected_draft_p"] = selq
        cases.append(ci)

    # ---- finalize_verification / resolve_stage_verification on fake caches --------------
    V = 128
    L = 12
    P = torch.softmax(gen_logits(g, (1, L + 1, V)), -1)
    Q = torch.softmax(gen_logits(g, (1, L, V)), -1)
    x = torch.randint(0, V, (1, L), generator=g)
    noise = torch.empty(1, V).exponential_(1, generator=g)
    out["finalize/P"] = P.numpy()
    out["finalize/Q"] = Q.numpy()
    out["finalize/x"] = x.numpy()
    out["finalize/noise"] = noise.numpy()
    for name, (prefix_len, gam, n) in {"reject": (8, 4, 9), "accept": (8, 4, 11)}.items():
        a = _FakeCache(Q.clone(), V)
        t = _FakeCache(P.clone(), V)
        with injected_noise([noise]):
            y = R.D.finalize_verification(a, t, x, prefix_len, gam, n)
        out[f"finalize/{name}/args"] = np.array([prefix_len, gam, n])
        out[f"finalize/{name}/out"] = y.numpy()
        out[f"finalize/{name}/rollbacks"] = np.array(a.rollback_calls + t.rollback_calls)

    prefix_len, gam = 8, 4
    for name, rvals in {"mixed": None, "all": 0.0}.items():
        a = _FakeCache(Q.clone(), V)
        t = _FakeCache(P.clone(), V)
        rr = torch.rand(1, gam, generator=g) if rvals is None else torch.full((1, gam), rvals)
        real_rand = torch.rand
        torch.rand = lambda *args, **kw: rr.clone()
        try:
            with injected_noise([noise]):
                ac, n, tk, alla = R.D.resolve_stage_verification(
                    a, t, x, prefix_len, gam, output_device=torch.device("cpu")
                )
        finally:
            torch.rand = real_rand
        out[f"resolve/{name}/r"] = rr.numpy()
        out[f"resolve/{name}/result"] = np.array([ac, n, int(tk), int(alla)])
        out[f"resolve/{name}/rollbacks"] = np.array(a.rollback_calls + t.rollback_calls)

    # ---- payload truncation -------------------------------------------------------------
    ph = torch.softmax(gen_logits(g, (2, 6, 50)), -1)
    xx = torch.randint(0, 50, (2, 9), generator=g)
    out["payload/ph"] = ph.numpy()
    out["payload/x"] = xx.numpy()
    for name, (pl, gm) in {"full": (3, 3), "rows": (5, 4), "toks": (8, 4), "none": (9, 2), "g0": (3, 0)}.items():
        tk, pq = R.D.collect_verification_payload(ph, xx, pl, gm)
        out[f"payload/{name}/args"] = np.array([pl, gm])
        out[f"payload/{name}/tokens"] = tk.numpy()
        out[f"payload/{name}/probs"] = pq.numpy()

    path = os.path.join(OUT, "hotpath_v1.npz")
    np.savez_compressed(path, **out)
    print(f"wrote {path}: {len(out)} arrays, {os.path.getsize(path) / 1e6:.2f} MB")


if __name__ == "__main__":
    main()
