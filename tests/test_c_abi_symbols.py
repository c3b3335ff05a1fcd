"""The C-ABI library loads without a GPU and exports every symbol include/duodec.h declares."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "duodec.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(duodec_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported_and_bound():
    from duodecoding_b200 import _lib
    names = _declared()
    assert len(names) >= 30
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in duodec.h but not exported by libduodec.so"
        assert n in _lib.SIGNATURES, f"{n} has no ctypes signature in duodecoding_b200/_lib.py"
    assert sorted(_lib.SIGNATURES) == names  # nothing bound that the header does not declare
    assert "sm_100a" in _lib.version()


def test_argument_errors_come_back_as_codes_without_a_gpu():
    from duodecoding_b200 import _lib
    lib = _lib.lib
    assert lib.duodec_norm_logits(None, 0, 0, None, 0, 0, 4, 0, 1.0, 0, 0.0, None) == -1      # V <= 0
    assert lib.duodec_norm_logits(None, 0, 0, None, 0, 0, 0, 16, 1.0, 0, 0.0, None) == 0      # no rows: no-op
    assert lib.duodec_shard_finish(None, 0, 4, 1, None, None, None) == -1                           # world <= 0
    assert lib.duodec_verify_workspace_bytes(64, 4) > 0 and lib.duodec_team_size(32000, 2) >= 1
    # the no-allocation variant refuses shapes whose scratch exceeds its built-in array before touching the device
    assert lib.duodec_verify_static_max_bytes() == 12 << 20
    assert lib.duodec_verify_workspace_bytes(64, 4) <= lib.duodec_verify_static_max_bytes()
    big = [None, None, 0, 0, 0, 0, 0, 100000, 8, 32000, 32000, None, 0, None, 0, None, 0, 1.0, 0, 0.0, 1.0, 0, 0.0, 0]
    assert lib.duodec_verify_workspace_bytes(100000, 8) > lib.duodec_verify_static_max_bytes()
    assert lib.duodec_verify_fused_static(*big, None, None, None, None, None, None, None) == -4


def test_kernels_are_registered_torch_library_ops():
    """Every kernel launch is a `duodec::` custom op whose outputs are annotated as mutated."""
    import torch

    import duodecoding_b200.ops  # noqa: F401  (registration happens at import)

    want = {"norm_logits": 1, "verify_fused": 7, "accept_probs": 4, "sample_rows": 1, "residual_sample": 1,
            "max_fn": 1, "topk_rebuild": 1, "kv_truncate": 1}
    for name, nmut in want.items():
        schema = getattr(torch.ops.duodec, name).default._schema
        mutated = [a for a in schema.arguments if a.alias_info is not None and a.alias_info.is_write]
        assert len(mutated) == nmut, (name, str(schema))
        assert len(schema.returns) == 0
    # the Meta implementation lets shape-only tracing pass through without a device
    x = torch.empty(3, 16, device="meta")
    torch.ops.duodec.max_fn(x, torch.empty(3, 16, device="meta"))
