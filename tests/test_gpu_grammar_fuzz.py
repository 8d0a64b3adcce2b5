"""a2' at scale: 60 000 random quantity strings through the CUDA parser, checked against the interpreter's own
`int()` (the authority the reference relies on, check-gpu-node.py:192) and against the C oracle."""
import random

import numpy as np
import pytest
import torch

from k8s_gpu_node_checker_b200 import engine, pnl
from oracle import pnl_oracle as po
from tests.helpers import F_DEFERRED, F_GPU, F_KEY0, KEYS, assert_same_outputs

pytestmark = pytest.mark.gpu

ALPHABET = list("0123456789") * 3 + list("_+- \t\n\x0b\x0c\r\x1c\x00kKi.e") + \
    ["٤", "４", "\xa0", " ", "\x85", "\U0001d7dc", "²", "\ud800", "​", "१", "๓"]


def _py(s):
    try:
        return int(s)
    except ValueError:
        return None


def test_random_quantities():
    rng = random.Random(20251106)
    values, rows = [], []
    for i in range(60000):
        n = rng.choice([0, 1, 1, 2, 3, 4, 6, 9, 14, 19, 20, 25])
        s = "".join(rng.choice(ALPHABET) for _ in range(n))
        if rng.random() < 0.3:
            s = rng.choice(["", "-", "+", " "]) + str(rng.randrange(0, 10 ** rng.choice([1, 3, 9, 17, 18, 19, 20, 24]))) + rng.choice(["", " ", "\n"])
        k = i % 4
        values.append((k, s))
        rows.append(pnl.encode_row([("cpu", "8"), (KEYS[k], s), ("pods", "110")], [("Ready", "True")]))
    p = pnl.from_rows(rows)
    clf = engine.Classifier(0)
    try:
        out = engine.DeviceResult.alloc(p.n_nodes, clf.device, defer_cap=p.n_nodes)
        clf.classify(engine.DeviceBatch.from_packed(p, clf.device), out)
        torch.cuda.synchronize()
        counts = out.counts.cpu().numpy()
        got = {"flags": out.flags.cpu().numpy(), "gpu_idx": out.gpu_idx.cpu().numpy().view(np.uint32),
               "gpu_total": out.gpu_total.cpu().numpy(), "gpu_vals": out.gpu_vals.cpu().numpy(), "counts": counts}
    finally:
        clf.close()
    assert_same_outputs(got, po.classify(p.rows, p.row_ptr, p.n_nodes, defer_cap=p.n_nodes), "fuzz vs C oracle")
    g = int(counts[0])
    pos_of = {int(n): i for i, n in enumerate(got["gpu_idx"][:g])}
    seen = {"kept": 0, "dropped": 0, "big": 0, "gpu": 0}
    for i, (k, s) in enumerate(values):
        want = _py(s) if s != "" else None                 # '' is falsy: skipped before int() (:188)
        f = int(got["flags"][i])
        if want is None:
            assert not f & (F_KEY0 << k) and not f & (F_GPU | F_DEFERRED), repr(s)
            seen["dropped"] += 1
        elif abs(want) >= 10 ** 18:
            assert f & F_DEFERRED and f & (F_KEY0 << k), repr(s)
            seen["big"] += 1
        else:
            assert f & (F_KEY0 << k) and not f & F_DEFERRED, repr(s)
            assert bool(f & F_GPU) == (want > 0), repr(s)
            seen["kept"] += 1
            if want > 0:
                assert int(got["gpu_vals"][pos_of[i], k]) == want and int(got["gpu_total"][pos_of[i]]) == want, repr(s)
                seen["gpu"] += 1
    assert min(seen.values()) > 500, seen
