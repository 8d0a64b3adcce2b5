"""Shared test helpers: expected device outputs from per-template truth."""
from __future__ import annotations

import hashlib
import json
from typing import Dict, List, Sequence

import numpy as np

KEYS = ("nvidia.com/gpu", "amd.com/gpu", "gpu.intel.com/i915", "intel.com/gpu")
F_GPU, F_READY, F_DEFERRED, F_KEY0, F_MALFORMED = 1, 2, 4, 8, 128
BIG = 10 ** 18


def spec_digest(spec) -> str:
    return hashlib.sha256(json.dumps(spec, sort_keys=True, ensure_ascii=True).encode()).hexdigest()[:16]


def template_truth(records: Sequence[Dict]) -> Dict[str, np.ndarray]:
    """Per-template flags / values from records {ready, gpus, bd} (golden or oracle made)."""
    n = len(records)
    flags = np.zeros(n, np.uint8)
    vals = np.zeros((n, 4), np.int64)
    total = np.zeros(n, np.int64)
    for i, r in enumerate(records):
        f = F_READY if r["ready"] else 0
        big = False
        for k, key in enumerate(KEYS):
            if key in r["bd"]:
                f |= F_KEY0 << k
                v = r["bd"][key]
                if abs(v) >= BIG:
                    big = True
                else:
                    vals[i, k] = v
        if big:
            f |= F_DEFERRED
        else:
            total[i] = int(vals[i].sum())
            assert total[i] == r["gpus"]
            if r["gpus"] > 0:
                f |= F_GPU
        flags[i] = f
    return {"flags": flags, "vals": vals, "total": total}


def expected_outputs(truth: Dict[str, np.ndarray], tid: np.ndarray, node_base: int = 0) -> Dict[str, np.ndarray]:
    """What the device (and the C oracle) must produce for a fleet drawn from the templates."""
    tid = np.asarray(tid, dtype=np.int64)
    flags = truth["flags"][tid]
    sel = np.nonzero(flags & F_GPU)[0]
    vals = truth["vals"][tid[sel]]
    counts = np.zeros(16, np.int64)
    counts[0] = len(sel)
    counts[1] = int(((flags[sel] & F_READY) != 0).sum())
    counts[2] = int(((flags & F_DEFERRED) != 0).sum())
    counts[3] = len(tid)
    for k in range(4):
        col = vals[:, k]
        if len(col) == 0 or int(np.abs(col).max()) < (1 << 36):
            t = int(col.sum())
        else:
            t = int(col.astype(object).sum())
        counts[4 + k] = np.array([t & ((1 << 64) - 1)], dtype=np.uint64).view(np.int64)[0]
        counts[8 + k] = t >> 64
    return {"flags": flags, "gpu_idx": (sel + node_base).astype(np.uint32), "gpu_total": truth["total"][tid[sel]],
            "gpu_vals": vals, "counts": counts, "deferred_nodes": (np.nonzero(flags & F_DEFERRED)[0] + node_base)}


def key_total(counts: np.ndarray, k: int) -> int:
    """Exact per-key total from the lo/hi pair of the counts block."""
    lo = int(counts[4 + k]) & ((1 << 64) - 1)
    return (int(counts[8 + k]) << 64) + lo


def assert_same_outputs(got: Dict, exp: Dict, what: str = ""):
    g = int(exp["counts"][0])
    assert int(got["counts"][0]) == g, f"{what}: GPU-node count {int(got['counts'][0])} != {g}"
    np.testing.assert_array_equal(got["counts"][:13], exp["counts"][:13], err_msg=f"{what}: counts")
    np.testing.assert_array_equal(got["flags"], exp["flags"], err_msg=f"{what}: flags")
    np.testing.assert_array_equal(np.asarray(got["gpu_idx"])[:g], exp["gpu_idx"], err_msg=f"{what}: gpu_idx")
    np.testing.assert_array_equal(np.asarray(got["gpu_total"])[:g], exp["gpu_total"], err_msg=f"{what}: gpu_total")
    np.testing.assert_array_equal(np.asarray(got["gpu_vals"]).reshape(-1, 4)[:g], exp["gpu_vals"],
                                  err_msg=f"{what}: gpu_vals")
