"""Drop-in for the reference's classification seam.

`list_gpu_nodes(api)` has the signature and return value of check-gpu-node.py:215-226: two lists of
six-key dicts (`name, ready, gpus, gpu_breakdown, labels, taints`, :202-212), GPU nodes in API order,
the Ready list holding the same dict objects.  The per-node decisions come from the B200 kernels;
this module only packs the nodes and dresses the winners with their pass-through payload.
"""
from __future__ import annotations

import json
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import engine, pnl

_default: Optional[engine.Classifier] = None


def default_classifier() -> engine.Classifier:
    global _default
    if _default is None:
        if not torch.cuda.is_available():
            raise RuntimeError("no CUDA device: the classification path runs on the GPU only (no CPU fallback)")
        _default = engine.Classifier(torch.cuda.current_device())
    return _default


def classify_packed(p: pnl.PackedNodeList, clf: Optional[engine.Classifier] = None) -> engine.FleetResult:
    """Host-resident PNL-v2 -> FleetResult through the host-buffer entry point (gnc_classify_host)."""
    clf = clf or default_classifier()
    out = clf.classify_host(p)
    return engine.finish(out["flags"][:p.n_nodes], out["gpu_idx"], out["gpu_total"], out["gpu_vals"], out["counts"],
                         out["defer"], p.rows)


def _payload(node: Any) -> Tuple[str, Dict, List]:
    """name / labels / taints exactly as check-gpu-node.py:203, :207-211 builds them."""
    md, spec = node.metadata, node.spec
    taints = [{"key": t.key, "value": t.value, "effect": t.effect} for t in (spec.taints or [])] \
        if spec and getattr(spec, "taints", None) else []
    return (md.name if md else ""), (md.labels if md and md.labels else {}), taints


def _native_args(res: engine.FleetResult):
    """FleetResult -> the buffers csrc/fastpack.c reads (bignum nodes travel in `wide`)."""
    wide = {n: (sum(bd.values()), bd) for n, bd in res.wide.items()} if res.wide else None
    totals = np.array([t if -2 ** 63 <= t < 2 ** 63 else 0 for t in res.gpu_total], dtype=np.int64) \
        if res.wide else np.asarray(res.gpu_total, dtype=np.int64)
    return (np.ascontiguousarray(res.gpu_idx, dtype=np.int64), np.ascontiguousarray(res.flags), totals,
            np.ascontiguousarray(res.gpu_vals, dtype=np.int64).reshape(-1), wide, res.node_base)


def build_records(nodes: Sequence[Any], res: engine.FleetResult, use_c: Optional[bool] = None) -> Tuple[List[Dict], List[Dict]]:
    """(gpu_nodes, ready_gpu_nodes) from the device's decisions plus each node's payload
    (check-gpu-node.py:202-212, :222-225).  The C extension builds the dicts when it is there."""
    fp = pnl._fastpack() if use_c is not False else None
    if fp is not None and hasattr(fp, "build_records"):
        return fp.build_records(nodes if isinstance(nodes, list) else list(nodes), *_native_args(res))
    gpu_nodes: List[Dict] = []
    ready_nodes: List[Dict] = []
    for pos, i in enumerate(res.gpu_idx.tolist()):
        name, labels, taints = _payload(nodes[i])
        rec = {"name": name, "ready": res.is_ready(pos), "gpus": res.gpu_total[pos],
               "gpu_breakdown": res.breakdown(pos), "labels": labels, "taints": taints}
        gpu_nodes.append(rec)
        if rec["ready"]:
            ready_nodes.append(rec)
    return gpu_nodes, ready_nodes


def classify_for_summary(nodes: Optional[Sequence[Any]], clf: Optional[engine.Classifier] = None):
    """(nodes as a list, FleetResult): what `summary_json_from_result` and `build_records` take."""
    nodes = list(nodes or [])
    return nodes, classify_packed(pnl.pack_nodes(nodes), clf)


class LastCall:
    """What the most recent `list_gpu_nodes` classified: lets the CLI wrapper write the `--json` document
    natively from the device's decisions instead of walking the list of dicts again (cli.py)."""
    nodes: Optional[List[Any]] = None
    result: Optional[engine.FleetResult] = None
    gpu_nodes: Optional[List[Dict]] = None


def classify_nodes(nodes: Optional[Sequence[Any]], clf: Optional[engine.Classifier] = None):
    nodes, res = classify_for_summary(nodes, clf)
    out = build_records(nodes, res)
    LastCall.nodes, LastCall.result, LastCall.gpu_nodes = nodes, res, out[0]
    return out


def list_gpu_nodes(api, clf: Optional[engine.Classifier] = None) -> Tuple[List[Dict], List[Dict]]:
    """Returns (gpu_nodes, ready_gpu_nodes) as list of dicts -- check-gpu-node.py:215-226."""
    return classify_nodes(api.list_node().items or [], clf)


def summary_json(gpu_nodes: List[Dict], ready_gpu_nodes: List[Dict]) -> str:
    """The `--json` document (check-gpu-node.py:274-279)."""
    return json.dumps({"total_nodes": len(gpu_nodes), "ready_nodes": len(ready_gpu_nodes), "nodes": gpu_nodes},
                      ensure_ascii=False, indent=2)


def summary_json_from_result(nodes: Sequence[Any], res: engine.FleetResult) -> str:
    """The same `--json` text straight from the device's decisions (check-gpu-node.py:273-279), written
    by the C emitter without building the list of dicts; falls back to build_records + json.dumps when
    the extension is not built or a payload value has a type it does not format."""
    fp = pnl._fastpack()
    if fp is not None:
        txt = fp.emit_summary_json(nodes, *_native_args(res))
        if txt is not None:
            return txt
    g, r = build_records(nodes, res)
    return summary_json(g, r)


def exit_code(gpu_nodes: List[Dict], ready_gpu_nodes: List[Dict]) -> int:
    """check-gpu-node.py:289-293."""
    if ready_gpu_nodes:
        return 0
    return 3 if gpu_nodes else 2
