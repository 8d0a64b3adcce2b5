"""Config 5 timing: raw NodeList JSON (1M nodes) -> GPU structural scan -> classify.  GPU box only.
    python tools/bench_json.py [nodes]"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from k8s_gpu_node_checker_b200 import engine, synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
tmpl = synth.g1_templates()
_, tid = synth.fleet("g1", 0, 0, n)
cache = {}
parts = []
for t in tid:
    t = int(t)
    if t not in cache:
        cache[t] = synth.node_json(tmpl[t], synth.g1_payload(0, tmpl[t]))
    parts.append(cache[t])
doc = ('{"kind":"NodeList","apiVersion":"v1","metadata":{"resourceVersion":"1"},"items":[' + ",".join(parts) + "]}").encode()
t0 = time.perf_counter()
root = json.loads(doc)
t_loads = time.perf_counter() - t0
del root
dev = torch.device("cuda", 0)
d = torch.frombuffer(bytearray(doc), dtype=torch.uint8).to(dev)
clf = engine.Classifier(0)
for _ in range(2):
    batch, info = clf.json_to_pnl(d, n)
out = engine.DeviceResult.alloc(n, dev)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
reps = 5
scan_ms, cls_ms = [], []
for _ in range(reps):
    ev[0].record()
    batch, info = clf.json_to_pnl(d, n)
    ev[1].record()
    clf.classify(batch, out)
    ev[2].record()
    torch.cuda.synchronize()
    scan_ms.append(ev[0].elapsed_time(ev[1]))
    cls_ms.append(ev[1].elapsed_time(ev[2]))
s, c = float(np.median(scan_ms)), float(np.median(cls_ms))
print(json.dumps({"nodes": n, "json_bytes": len(doc), "json_bytes_per_node": len(doc) / n, "scan_ms": s, "classify_ms": c,
                  "scan_GBps": len(doc) / s / 1e6, "nodes_per_s_scan_plus_classify": n / ((s + c) * 1e-3),
                  "pnl_rows_bytes": batch.rows_bytes, "general_parser_nodes": int(info[5]), "gpu_nodes": int(out.counts[0]),
                  "host_json_loads_s": t_loads, "host_json_loads_MBps": len(doc) / t_loads / 1e6}))
