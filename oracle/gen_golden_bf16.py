"""Generate tests/golden/bf16_v1.npz with the UNMODIFIED reference IN BFLOAT16 (run in the build container only).

The reference loads its models in bf16 (src/engine.py:261) and computes ``logits / T``, the softmax, the ratio
``p / q`` and the comparison with ``r`` in that dtype (src/utils.py:686-709, src/decoding_ops.py:127-147).  The
kernels read the same bf16 logits and compute in fp32.  The fixture holds the reference's bf16 results so that the
GPU tests can state the canonicalisation exactly: decisions agree wherever ``|p/q - r|`` is outside bf16 rounding."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
import src.decoding_ops as D  # noqa: E402
import src.decoding_types as T  # noqa: E402
import src.utils as U  # noqa: E402


def main():
    g = torch.Generator().manual_seed(909)
    out = {}
    for ci, (Bn, gamma, V, sigma, temp, top_k, top_p) in enumerate([(16, 4, 2000, 0.7, 1.0, 0, 0.0),
                                                                     (16, 4, 2000, 1.0, 1.0, 0, 0.9),
                                                                     (8, 6, 4096, 0.5, 0.7, 50, 0.0)]):
        tl = (torch.randn(Bn, gamma + 1, V, generator=g) * 3).bfloat16()
        dl = (tl[:, :gamma].float() + sigma * torch.randn(Bn, gamma, V, generator=g)).bfloat16()
        q = U.norm_logits(dl, temp, top_k, top_p)          # bf16 in, bf16 out
        p = U.norm_logits(tl, temp, top_k, top_p)
        assert q.dtype == torch.bfloat16 and p.dtype == torch.bfloat16
        toks = torch.stack([torch.multinomial(q[:, j].float().clamp_min(0), 1, generator=g)[:, 0] for j in range(gamma)], 1)
        r = torch.rand(Bn, gamma, generator=g)
        masks, sq, sp, counts = [], [], [], []
        for b in range(Bn):  # the reference reads row 0: one call per sequence
            vi = T.VerificationInputs(draft_probs_batch=q[b:b + 1], target_probs_batch=p[b:b + 1, :gamma],
                                      draft_tokens=toks[b:b + 1], draft_token_indices=toks[b:b + 1].unsqueeze(-1),
                                      prefix_len=1, gamma=gamma, actual_gamma=gamma, max_idx=gamma)
            res = D.compute_acceptance_result(vi, r=r[b:b + 1])
            masks.append(res.accept_mask[0]), counts.append(res.accepted_count)
            sq.append(res.selected_draft_p[0].float()), sp.append(res.selected_target_p[0].float())
        pre = f"bf16/{ci}/"
        out[pre + "draft"] = dl.view(torch.int16).numpy()      # the bf16 bit patterns
        out[pre + "target"] = tl.view(torch.int16).numpy()
        out[pre + "tokens"], out[pre + "r"] = toks.numpy(), r.numpy()
        out[pre + "params"] = np.array([temp, top_k, top_p], dtype=np.float64)
        out[pre + "accept_mask"] = torch.stack(masks).numpy()
        out[pre + "accepted_count"] = np.array(counts, dtype=np.int64)
        out[pre + "selected_draft_p"], out[pre + "selected_target_p"] = torch.stack(sq).numpy(), torch.stack(sp).numpy()
    path = os.path.join(ROOT, "tests", "golden", "bf16_v1.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KB;", {k: float(v.mean()) for k, v in out.items() if k.endswith("accepted_count")})


if __name__ == "__main__":
    main()
