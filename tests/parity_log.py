"""Makes the tolerance gates of the parity tests observable (VERDICT r01 "What's weak" #2).

Every gradient / image comparison appends one JSON record: which test, which quantity, the achieved
relative L2 error against the fp32 oracle, whether the STRICT gate (<= tol against the fp32 oracle)
passed or the comparison went through a RELAXED branch (and which), the fp32-oracle noise against the
fp64 oracle, and the fraction of rays masked as face-sensitive. ``tools/parity_report.py`` turns the
records into ``profiles/rNN_parity.md``.

The log goes to ``$DVR_PARITY_LOG`` or ``gpurun_out/parity_records.jsonl`` (``gpurun_out/`` is what
travels back from the GPU box); failures to write are ignored -- the log never decides a test.
"""
from __future__ import annotations

import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_current_test = {"id": "?"}


def set_current_test(nodeid: str):
    _current_test["id"] = nodeid


def path() -> str:
    return os.environ.get("DVR_PARITY_LOG", os.path.join(ROOT, "gpurun_out", "parity_records.jsonl"))


def record(quantity: str, gate: str, **fields):
    """gate: "strict" | "relaxed:<branch>" | "info"."""
    rec = {"test": _current_test["id"], "quantity": quantity, "gate": gate}
    for k, v in fields.items():
        rec[k] = float(v) if hasattr(v, "__float__") and not isinstance(v, (bool, str)) else v
    try:
        p = path()
        os.makedirs(os.path.dirname(p), exist_ok=True)
        with open(p, "a") as f:
            f.write(json.dumps(rec) + "\n")
    except OSError:
        pass
    return rec
