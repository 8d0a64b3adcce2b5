"""Raw /api/v1/nodes JSON -> the objects the reference reads (TEST INFRASTRUCTURE).

The reference gets its nodes from `client.CoreV1Api().list_node().items` (check-gpu-node.py:217,
:253); the JSON -> object step lives in the third-party `kubernetes` client (33.1.0, uv.lock:104-105),
whose source is not under /root/reference.  **Parity at this boundary is unpinned**: the reference has
no tests or fixtures for it.  This adapter restates the published behaviour that matters to
the path -- `json.loads` (duplicate object keys: last wins; `\\uXXXX` escapes, surrogate pairs joined) and
the client's `dict(str, str)` deserialisation of `status.capacity` (`str(value)` for non-strings, `None`
kept) -- and then runs the reference's own functions (through oracle/cgn_oracle.py) on the result.
"""
from __future__ import annotations

import json
from typing import Any, List

from . import ensure_kubernetes_importable

ensure_kubernetes_importable()
from kubernetes import client as kc  # noqa: E402


def _capacity(raw: Any):
    if not isinstance(raw, dict):
        return None
    out = {}
    for k, v in raw.items():
        if v is None or isinstance(v, str):
            out[k] = v
        elif isinstance(v, bool) or isinstance(v, (int, float)):
            out[k] = v if not v else str(v)      # falsy numbers stay falsy (skipped, :188); others as the client's str()
        else:
            out[k] = None                         # nested containers never survive int(str(.)): absent either way
    return out


def node_from_json(item: Any):
    """One element of `items` -> stub V1Node with the attribute names check-gpu-node.py reads."""
    if not isinstance(item, dict):
        return kc.V1Node()
    md, spec, st = item.get("metadata"), item.get("spec"), item.get("status")
    status = None
    if isinstance(st, dict):
        conds = None
        if isinstance(st.get("conditions"), list):
            conds = [kc.V1NodeCondition(type=c.get("type"), status=c.get("status"))
                     for c in st["conditions"] if isinstance(c, dict)]
        status = kc.V1NodeStatus(capacity=_capacity(st.get("capacity")), conditions=conds)
    meta = None
    if isinstance(md, dict):
        meta = kc.V1ObjectMeta(name=md.get("name"), labels=md.get("labels"))
    sp = None
    if isinstance(spec, dict):
        taints = spec.get("taints")
        sp = kc.V1NodeSpec(taints=[kc.V1Taint(key=t.get("key"), value=t.get("value"), effect=t.get("effect"))
                                   for t in taints if isinstance(t, dict)] if isinstance(taints, list) else None)
    return kc.V1Node(metadata=meta, spec=sp, status=status)


def nodes_from_json(doc: bytes) -> List[Any]:
    """`api.list_node().items` for a NodeList document."""
    root = json.loads(doc)
    items = root.get("items") if isinstance(root, dict) else None
    return [node_from_json(it) for it in items] if isinstance(items, list) else []
