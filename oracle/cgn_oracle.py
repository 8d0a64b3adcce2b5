"""Python restatement of the reference's classification path (TEST INFRASTRUCTURE).

Follows /root/reference/check-gpu-node.py (`cgn.py` below); every function cites the lines it
restates.  Works on objects with the kubernetes-client attribute names (the `k8s_stub` models or
the real client).  Pinned against the reference itself by tests/golden (see oracle/__init__.py).

Also holds `int_grammar`, an explicit spelling of CPython's `int(str)` base-10 grammar, which is
what the C oracle and the CUDA kernel implement; tests pin it to the interpreter's own `int`.
"""
from __future__ import annotations

import json
import sys
import unicodedata
from typing import Any, Dict, List, Optional, Tuple

from . import ensure_kubernetes_importable

ensure_kubernetes_importable()
from kubernetes.client import V1NodeCondition  # noqa: E402

# cgn.py:39-44 -- list order is the key order of every `gpu_breakdown`.
GPU_KEYS: Tuple[str, ...] = ("nvidia.com/gpu", "amd.com/gpu", "gpu.intel.com/i915", "intel.com/gpu")


def ready_of(node: Any) -> bool:
    """cgn.py:172-178: any V1NodeCondition with type 'Ready' and status 'True'."""
    st = node.status
    if not st or not st.conditions:                       # :173-174
        return False
    return any(                                            # :175-177 (first hit returns)
        isinstance(c, V1NodeCondition) and c.type == "Ready" and c.status == "True"
        for c in st.conditions
    )


def gpu_breakdown_of(node: Any) -> Dict[str, int]:
    """cgn.py:181-196: capacity only (allocatable never read); falsy skipped; int(str(v)) or drop."""
    out: Dict[str, int] = {}
    st = node.status or None                               # :183
    if not st or not st.capacity:                          # :184
        return out
    for key in GPU_KEYS:                                   # :186
        raw = st.capacity.get(key)                         # :187
        if not raw:                                        # :188  ("0" is truthy and stays)
            continue
        try:
            out[key] = int(str(raw))                       # :192
        except Exception:                                  # :193-195
            pass
    return out


def node_record(node: Any) -> Dict[str, Any]:
    """cgn.py:199-212: the six-key record, keys in this order."""
    caps = gpu_breakdown_of(node)                          # :200
    md, spec = node.metadata, node.spec
    return {
        "name": md.name if md else "",                                         # :203
        "ready": ready_of(node),                                               # :204
        "gpus": sum(caps.values()) if caps else 0,                             # :201, :205
        "gpu_breakdown": caps,                                                 # :206
        "labels": md.labels if md and md.labels else {},                       # :207
        "taints": [{"key": t.key, "value": t.value, "effect": t.effect}        # :208-211
                   for t in (spec.taints or [])] if spec and getattr(spec, "taints", None) else [],
    }


def classify_fleet(nodes) -> Tuple[List[Dict], List[Dict]]:
    """cgn.py:215-226 minus the API call: keep gpus > 0 in order; Ready subset shares the dicts."""
    gpu_nodes: List[Dict] = []
    ready_nodes: List[Dict] = []
    for n in nodes or []:                                  # :217 `or []`, :220
        rec = node_record(n)                               # :221
        if rec["gpus"] > 0:                                # :222
            gpu_nodes.append(rec)                          # :223
            if rec["ready"]:                               # :224
                ready_nodes.append(rec)                    # :225
    return gpu_nodes, ready_nodes


def list_gpu_nodes(api) -> Tuple[List[Dict], List[Dict]]:
    """cgn.py:215-226 with the same signature as the reference."""
    return classify_fleet(api.list_node().items or [])


def summary_payload(gpu_nodes: List[Dict], ready_nodes: List[Dict]) -> Dict[str, Any]:
    """cgn.py:274-278."""
    return {"total_nodes": len(gpu_nodes), "ready_nodes": len(ready_nodes), "nodes": gpu_nodes}


def summary_json(gpu_nodes: List[Dict], ready_nodes: List[Dict]) -> str:
    """cgn.py:279 -- the exact text `--json` prints."""
    return json.dumps(summary_payload(gpu_nodes, ready_nodes), ensure_ascii=False, indent=2)


def exit_code(gpu_nodes: List[Dict], ready_nodes: List[Dict]) -> int:
    """cgn.py:289-293."""
    if ready_nodes:
        return 0
    if gpu_nodes:
        return 3
    return 2


def key_totals(gpu_nodes: List[Dict]) -> List[int]:
    """Fleet-wide per-key totals over the GPU nodes (the north star's extra reduction)."""
    return [sum(n["gpu_breakdown"].get(k, 0) for n in gpu_nodes) for k in GPU_KEYS]


# --------------------------------------------------------------------------------------------
# int(str) grammar, spelled out (cgn.py:192 relies on the interpreter for this).
# CPython: PyLong_FromUnicodeObject -> _PyUnicode_TransformDecimalAndSpaceToASCII ->
# PyLong_FromString(base 10).  Code points < 127 pass through; >= 127 become ' ' if
# Py_UNICODE_ISSPACE, an ASCII digit if they have a decimal value, else the parse fails.
# Then: [ \t\n\v\f\r]* [+-]? digit ('_'? digit)* [ \t\n\v\f\r]* and nothing else; more than
# sys.get_int_max_str_digits() digit characters (leading zeros count, '_' do not) fails.
# --------------------------------------------------------------------------------------------
_ASCII_WS = frozenset(b"\t\n\v\f\r ")


def _wide_space(cp: int) -> bool:
    return cp >= 127 and chr(cp).isspace()


def _wide_digit(cp: int) -> int:
    return unicodedata.decimal(chr(cp), -1) if cp >= 127 else -1


def int_grammar(text: str, max_digits: Optional[int] = None) -> Optional[int]:
    """Value of `int(text)` or None where `int` raises.  No call to `int` on the text."""
    if max_digits is None:
        max_digits = sys.get_int_max_str_digits()
    sym: List[int] = []          # ASCII code per character after the transform
    for ch in text:
        cp = ord(ch)
        if cp < 127:
            sym.append(cp)
        elif _wide_space(cp):
            sym.append(0x20)
        else:
            d = _wide_digit(cp)
            if d < 0:
                return None
            sym.append(0x30 + d)
    i, n = 0, len(sym)
    while i < n and sym[i] in _ASCII_WS:
        i += 1
    neg = False
    if i < n and sym[i] in (0x2B, 0x2D):
        neg = sym[i] == 0x2D
        i += 1
    value, digits, prev_us = 0, 0, True      # prev_us=True forbids a leading '_'
    while i < n:
        c = sym[i]
        if 0x30 <= c <= 0x39:
            value = value * 10 + (c - 0x30)
            digits += 1
            prev_us = False
        elif c == 0x5F:
            if prev_us:
                return None
            prev_us = True
        else:
            break
        i += 1
    if digits == 0 or prev_us:
        return None
    while i < n and sym[i] in _ASCII_WS:
        i += 1
    if i != n:
        return None
    if max_digits and digits > max_digits:
        return None
    return -value if neg else value


def unicode_decimal_runs() -> List[int]:
    """Start code points (>= 127) of every run of ten consecutive decimal digits 0..9."""
    starts = []
    cp = 127
    while cp < 0x110000:
        if unicodedata.decimal(chr(cp), -1) == 0:
            assert all(unicodedata.decimal(chr(cp + d), -1) == d for d in range(10)), hex(cp)
            starts.append(cp)
            cp += 10
        else:
            assert unicodedata.decimal(chr(cp), -1) == -1, hex(cp)
            cp += 1
    return starts


def unicode_wide_spaces() -> List[int]:
    """Code points >= 127 that `int` treats as whitespace."""
    return [cp for cp in range(127, 0x110000) if chr(cp).isspace()]
