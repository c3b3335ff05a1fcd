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
    assert len(names) >= 18
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
