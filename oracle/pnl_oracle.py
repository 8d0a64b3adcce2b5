"""ctypes binding + build recipe for oracle/pnl_oracle.c (TEST INFRASTRUCTURE)."""
from __future__ import annotations

import ctypes
import os
import subprocess
import sys
from typing import Dict

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "libpnl_oracle.so")
SRC = os.path.join(HERE, "pnl_oracle.c")


class Defer(ctypes.Structure):
    _fields_ = [("node", ctypes.c_uint32), ("off", ctypes.c_uint32 * 4), ("len", ctypes.c_uint32 * 4)]


def build(force: bool = False) -> str:
    deps = [SRC, os.path.join(HERE, "pnl_unicode.h")]
    if force or not os.path.exists(SO) or any(os.path.getmtime(d) > os.path.getmtime(SO) for d in deps):
        subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-Wall", "-Wextra", "-o", SO, SRC])
    return SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
        _lib.pnl_classify.restype = ctypes.c_int
        _lib.pnl_flags_only.restype = ctypes.c_int
        _lib.pnl_parse_quantity.restype = ctypes.c_int
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def classify(rows: np.ndarray, row_ptr: np.ndarray, n: int, node_base: int = 0, max_digits: int | None = None,
             defer_cap: int = 1 << 16) -> Dict[str, np.ndarray]:
    """Run the C restatement; returns trimmed numpy outputs keyed like the device result."""
    if max_digits is None:
        max_digits = sys.get_int_max_str_digits()
    rows = np.ascontiguousarray(rows, dtype=np.uint8)
    row_ptr = np.ascontiguousarray(row_ptr).view(np.uint32)
    flags = np.zeros(n, np.uint8)
    idx = np.zeros(max(n, 1), np.uint32)
    tot = np.zeros(max(n, 1), np.int64)
    vals = np.zeros(max(n, 1) * 4, np.int64)
    counts = np.zeros(16, np.int64)
    defer = (Defer * max(defer_cap, 1))()
    rc = lib().pnl_classify(_p(rows), _p(row_ptr), ctypes.c_uint32(n), ctypes.c_uint32(node_base),
                            ctypes.c_int(max_digits), _p(flags), _p(idx), _p(tot), _p(vals), _p(counts),
                            ctypes.byref(defer), ctypes.c_uint32(defer_cap))
    assert rc == 0
    g, d = int(counts[0]), min(int(counts[2]), defer_cap)
    recs = [(defer[i].node, tuple(defer[i].off), tuple(defer[i].len)) for i in range(d)]
    return {"flags": flags, "gpu_idx": idx[:g].copy(), "gpu_total": tot[:g].copy(),
            "gpu_vals": vals[:g * 4].reshape(g, 4).copy(), "counts": counts, "defer": recs}


def parse_quantity(data: bytes, max_digits: int | None = None):
    """(status, value): status 0 absent, 1 ok, 2 big."""
    if max_digits is None:
        max_digits = sys.get_int_max_str_digits()
    out = ctypes.c_int64(0)
    buf = (ctypes.c_uint8 * max(len(data), 1)).from_buffer_copy(data or b"\0")
    st = lib().pnl_parse_quantity(buf, ctypes.c_uint32(len(data)), ctypes.c_int(max_digits), ctypes.byref(out))
    return st, out.value
