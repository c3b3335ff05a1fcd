"""Generate tests/golden/ranks_v1.npz with the UNMODIFIED reference (run in the build container only).

``src/baselines.py`` does not import here (it pulls in ``SpecDec_pp``), so the three rank helpers
(``_compute_token_vocab_rank``, ``_compute_transfer_topk_rank``, ``_record_accepted_token_ranks``, lines 121-168)
are taken from the reference's SOURCE TEXT at generation time: their function definitions are cut out of the file
with ``ast`` and executed unchanged.  Inputs are continuous random probabilities (no equal values, so torch.topk's
unspecified order among ties cannot enter the fixture)."""
import ast
import os
import sys
from typing import List, Optional  # noqa: F401  (names used by the extracted source)

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = "/root/reference/src/baselines.py"
NAMES = ("_compute_token_vocab_rank", "_compute_transfer_topk_rank", "_record_accepted_token_ranks")


def reference_functions():
    text = open(SRC).read()
    tree = ast.parse(text)
    ns = {"torch": torch, "List": List, "Optional": Optional}
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in NAMES:
            exec(compile(ast.Module(body=[node], type_ignores=[]), SRC, "exec"), ns)  # noqa: S102
    return [ns[n] for n in NAMES]


def main():
    _, _, record = reference_functions()
    g = torch.Generator().manual_seed(77)
    out = {}
    for ci, (V, L, prefix, accepted, k) in enumerate([(500, 12, 4, 5, 64), (2048, 9, 3, 4, 0), (777, 20, 10, 8, 16),
                                                      (300, 6, 2, 3, 299), (1000, 8, 5, 0, 32), (640, 10, 6, 4, 1)]):
        probs = torch.softmax(torch.randn(1, L, V, generator=g) * 2.5, -1)
        x = torch.randint(0, V, (1, L + 1), generator=g)
        for i in range(accepted):  # mostly likely tokens, as accepted tokens are
            if i % 3 != 2:
                x[0, prefix + i] = torch.multinomial(probs[0, prefix + i - 1], 1, generator=g)
        vr, it, tr = [], [], []
        record(stage_probs=probs, x=x, prefix_len=prefix, accepted_count=accepted, transfer_top_k=k,
               vocab_rank_history=vr, in_transfer_topk_history=it, transfer_topk_rank_history=tr)
        pre = f"ranks/{ci}/"
        out[pre + "probs"], out[pre + "x"] = probs.numpy(), x.numpy()
        out[pre + "params"] = np.array([prefix, accepted, k], dtype=np.int64)
        out[pre + "vocab_rank"] = np.array(vr, dtype=np.int64)
        out[pre + "in_topk"] = np.array(it, dtype=np.bool_)
        out[pre + "topk_rank"] = np.array(tr, dtype=np.int64)
    path = os.path.join(ROOT, "tests", "golden", "ranks_v1.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, {k: v.tolist() for k, v in out.items() if k.endswith("rank") or k.endswith("in_topk")})


if __name__ == "__main__":
    sys.exit(main())
