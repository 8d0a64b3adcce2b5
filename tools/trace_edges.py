"""Start-up and tail of the classify kernel from the in-kernel timestamps (debug build libgnc_trace.so, GPU box only).

    python tools/trace_edges.py [nodes]

Prints, relative to the first CTA's entry: when CTAs enter, when their first tile is in shared memory,
when they have walked their last tile, drained the record ring, and left; and the kernel time by CUDA events."""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from k8s_gpu_node_checker_b200 import native, engine, synth  # noqa: E402

native.LIB_PATH = os.path.join(native.CSRC, "libgnc_trace.so")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_250_000
tmpl = synth.g1_templates()
rows_t = synth.spec_rows(tmpl)
tid = synth.g1_tid(0, 0, n)
rows, rp, total = synth.expand(rows_t, torch.from_numpy(tid).cuda())
clf = engine.Classifier(0)
batch = engine.DeviceBatch(rows, rp, n, total)
out = engine.DeviceResult.alloc(n, clf.device)
for _ in range(3):
    clf.classify(batch, out)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ms = []
for _ in range(10):
    e0.record(); clf.classify(batch, out); e1.record(); torch.cuda.synchronize(); ms.append(e0.elapsed_time(e1))
trace = torch.zeros(296 * 64 * 8, dtype=torch.int64, device="cuda")
clf.lib.gnc_debug_trace.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
clf.lib.gnc_debug_trace(clf._h, ctypes.c_void_p(trace.data_ptr()))
clf.classify(batch, out)
torch.cuda.synchronize()
t = trace.cpu().numpy().reshape(296, 64, 8).astype(np.int64)
entry, last_walk, drained, leave = t[:, 0, 7], t[:, 1, 7], t[:, 2, 7], t[:, 3, 7]
first_full = t[:, 0, 0]
ok = entry > 0
z = entry[ok].min()
def show(name, x):
    x = x[ok & (x > 0)]
    print(f"{name:28s} min {(x.min() - z) / 1e3:7.2f}  med {(np.median(x) - z) / 1e3:7.2f}  max {(x.max() - z) / 1e3:7.2f} us")
print(f"nodes {n}, kernel by events: median {np.median(ms) * 1e3:.1f} us (trace build), CTAs traced {int(ok.sum())}")
show("CTA entry", entry)
show("first tile in shared memory", first_full)
show("last tile walked", last_walk)
show("ring drained (all warps)", drained)
show("CTA exit", leave)
