"""Generate tests/golden/cuhlm_v1.npz with the UNMODIFIED reference (run in the build container only):
``CUHLM.calculate_uncertainty`` with ``torch.rand`` / ``torch.multinomial`` replaced by injected draws, and
``_calculate_compressed_vocab_size`` on small vocabularies (the reference's search is O(V^2) Python)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import src.communication as C  # noqa: E402


def main():
    g = torch.Generator().manual_seed(2024)
    out = {}
    cls = [c for c in (getattr(C, n) for n in dir(C)) if isinstance(c, type) and hasattr(c, "_calculate_compressed_vocab_size")][0]
    # ---- calculate_uncertainty: inject temperatures (torch.rand * theta_max) and the multinomial noise
    for ci, (V, M, theta_max, scale) in enumerate([(300, 20, 2.0, 3.0), (2048, 20, 2.0, 1.5), (777, 8, 0.5, 4.0)]):
        logits = torch.randn(V, generator=g) * scale
        u = torch.rand(M, generator=g)
        noise = torch.empty(M, V).exponential_(1, generator=g)
        draft = int(torch.argmax(logits))
        real_rand, real_multi = torch.rand, torch.multinomial
        torch.rand = lambda *a, **k: u.clone()
        torch.multinomial = lambda p, n, *a, **k: torch.argmax(p / noise, dim=-1, keepdim=True)
        try:
            val = cls.calculate_uncertainty(logits, M=M, theta_max=theta_max, draft_token=draft)
        finally:
            torch.rand, torch.multinomial = real_rand, real_multi
        pre = f"unc/{ci}/"
        out[pre + "logits"], out[pre + "u"], out[pre + "noise"] = logits.numpy(), u.numpy(), noise.numpy()
        out[pre + "params"] = np.array([M, theta_max, draft], dtype=np.float64)
        out[pre + "value"] = np.array([val])
    # ---- _calculate_compressed_vocab_size
    for ci, (V, scale, unc, theta, use_draft) in enumerate([(300, 3.0, 0.3, 0.1, True), (512, 1.0, 0.8, 0.1, False),
                                                            (400, 5.0, 0.05, 0.02, True), (350, 0.2, 0.5, 0.1, True),
                                                            (256, 2.0, 1.0, 0.5, False)]):
        probs = torch.softmax(torch.randn(V, generator=g) * scale, -1)
        sim = cls(bandwidth_edge_cloud=100.0, vocab_size=V)
        draft = int(torch.multinomial(probs, 1, generator=g)) if use_draft else None
        k = sim._calculate_compressed_vocab_size(unc, probs, theta=theta, draft_token=draft)
        pre = f"kstar/{ci}/"
        out[pre + "probs"] = probs.numpy()
        out[pre + "params"] = np.array([unc, theta, -1 if draft is None else draft], dtype=np.float64)
        out[pre + "k"] = np.array([k])
    path = os.path.join(ROOT, "tests", "golden", "cuhlm_v1.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, len(out), "arrays;", {k: v.tolist() for k, v in out.items() if k.endswith("/k") or k.endswith("/value")})


if __name__ == "__main__":
    main()
