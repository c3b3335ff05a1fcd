"""Mount the B200 path over an importable copy of the reference (``src`` package) without editing it.

``install()`` rebinds the hot-path names inside ``src.utils``, ``src.decoding_ops``,
``src.proposal_utils``, ``src.model_gpu`` and the static helpers of
``src.communication.CommunicationSimulator`` to this package's implementations.  Only module globals
and class attributes are assigned, so late binding - and with it
``unittest.mock.patch("src.decoding_ops.sample")`` - keeps working; every registered
``eval_mode`` of ``src/register.py`` reaches the kernels through those names.  ``uninstall()`` restores
the originals.
"""

from __future__ import annotations

import importlib
from typing import Dict, List, Tuple

from . import baselines, communication, decoding_ops, model_gpu, proposal_utils, utils

_PLAN = {
    "utils": (utils, ["top_k_top_p_filter", "norm_logits", "sample", "max_fn", "rebuild_topk_probs",
                      "rebuild_topk_uniform_probs"]),
    "decoding_ops": (decoding_ops, ["collect_verification_payload", "prepare_verification_inputs",
                                    "compute_acceptance_result", "compute_residual_distribution",
                                    "build_rollback_plan", "apply_rollback", "sample_reject_token",
                                    "sample_accept_token", "verify_draft_sequence", "verify_draft_sequence_result",
                                    "resolve_stage_verification", "finalize_verification", "sample", "max_fn",
                                    "rebuild_topk_uniform_probs"]),
    "proposal_utils": (proposal_utils, ["proposal_top_k", "build_draft_probs_override", "stage_prob_history"]),
    "model_gpu": (model_gpu, ["KVCacheModel", "norm_logits", "sample", "rebuild_topk_uniform_probs"]),
    # (src.baselines imports optional packages: skipped when it does not import)
    "baselines": (baselines, ["_compute_token_vocab_rank", "_compute_transfer_topk_rank",
                              "_record_accepted_token_ranks"]),
}
_SAVED: List[Tuple[object, str, object]] = []


def install(package: str = "src") -> Dict[str, List[str]]:
    """Rebind the names; returns {module: [names rebound]}.  Modules that do not import are skipped."""
    done: Dict[str, List[str]] = {}
    for mod_name, (ours, names) in _PLAN.items():
        try:
            theirs = importlib.import_module(f"{package}.{mod_name}")
        except Exception:  # noqa: BLE001  (a reference module with missing optional dependencies)
            continue
        for name in names:
            if hasattr(theirs, name) and hasattr(ours, name):
                _SAVED.append((theirs, name, getattr(theirs, name)))
                setattr(theirs, name, getattr(ours, name))
                done.setdefault(mod_name, []).append(name)
    try:
        sim = importlib.import_module(f"{package}.communication").CommunicationSimulator
        for name in ("_apply_top_k_compression", "rebuild_full_probs", "compress_rebuild_probs"):
            _SAVED.append((sim, name, sim.__dict__[name]))
            setattr(sim, name, staticmethod(getattr(communication, name.lstrip("_"))))
            done.setdefault("communication", []).append(name)
        cuhlm = getattr(importlib.import_module(f"{package}.communication"), "CUHLM", None)
        if cuhlm is not None:  # the CUHLM helpers (src/communication.py:541-575, :681-755)
            _SAVED.append((cuhlm, "calculate_uncertainty", cuhlm.__dict__["calculate_uncertainty"]))
            cuhlm.calculate_uncertainty = staticmethod(communication.calculate_uncertainty)
            _SAVED.append((cuhlm, "_calculate_compressed_vocab_size", cuhlm.__dict__["_calculate_compressed_vocab_size"]))
            cuhlm._calculate_compressed_vocab_size = communication.CUHLM._calculate_compressed_vocab_size
            done.setdefault("communication", []).extend(["CUHLM.calculate_uncertainty",
                                                         "CUHLM._calculate_compressed_vocab_size"])
    except Exception:  # noqa: BLE001
        pass
    return done


def uninstall() -> None:
    while _SAVED:
        obj, name, old = _SAVED.pop()
        setattr(obj, name, old)
