"""Shared check of the bf16 reference fixtures (tests/golden/bf16_v1.npz, oracle/gen_golden_bf16.py): used with the
oracle on CPU and with the CUDA path on the GPU."""
import os

import numpy as np
import torch


def check_against_bf16_goldens(run):
    """run(draft bf16, target bf16, tokens, r, T, top_k, top_p) -> dict of CPU tensors (verify_fused's keys)."""
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "bf16_v1.npz"))
    cases = sorted({k.split("/")[1] for k in z.files})
    assert len(cases) == 3
    total = flips = 0
    for ci in cases:
        pre = f"bf16/{ci}/"
        T, top_k, top_p = (float(v) for v in z[pre + "params"])
        dl = torch.from_numpy(z[pre + "draft"]).view(torch.bfloat16)
        tl = torch.from_numpy(z[pre + "target"]).view(torch.bfloat16)
        toks, r = torch.from_numpy(z[pre + "tokens"]), torch.from_numpy(z[pre + "r"])
        got = run(dl, tl, toks, r, T, int(top_k), top_p)
        ref_mask = torch.from_numpy(z[pre + "accept_mask"])
        ref_q, ref_p = torch.from_numpy(z[pre + "selected_draft_p"]), torch.from_numpy(z[pre + "selected_target_p"])
        gq, gp = got["selected_draft_p"].float(), got["selected_target_p"].float()
        # a bf16 softmax carries 8 bits; with T != 1 the reference also rounds logits / T to bf16 (0.4 % of |x| ~ 10
        # in the exponent: a few per cent of the probability)
        tol = 1e-2 if T == 1.0 else 6e-2
        torch.testing.assert_close(gq, ref_q, rtol=tol, atol=1e-6)
        torch.testing.assert_close(gp, ref_p, rtol=tol, atol=1e-6)
        assert torch.equal(gq == 0, ref_q == 0) and torch.equal(gp == 0, ref_p == 0)  # the same filter supports
        ratio = (gp / gq).double()
        margin = (ratio - r.double()).abs() / r.double().clamp_min(1e-30)
        differ = got["accept_mask"] != ref_mask
        total += differ.numel()
        flips += int(differ.sum())
        assert bool((margin[differ] < 2.5 * tol).all()), (ci, margin[differ])  # only inside bf16 rounding of p/q
        # accepted counts: identical wherever no decision of the sequence sits inside that margin
        safe = (margin >= 2.5 * tol).all(1) | ~torch.isfinite(ratio).all(1)
        ref_n = torch.from_numpy(z[pre + "accepted_count"])
        assert torch.equal(got["accepted_len"][safe], ref_n[safe]), ci
        assert int(safe.sum()) >= len(safe) // 2
    assert flips <= 0.03 * total, (flips, total)
