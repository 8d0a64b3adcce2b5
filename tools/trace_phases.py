"""Per-iteration phase timeline of the classify kernel (debug build libgnc_trace.so, GPU box only).

    python tools/trace_phases.py [nodes]
"""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from k8s_gpu_node_checker_b200 import native, engine, synth  # noqa: E402

native.LIB_PATH = os.path.join(native.CSRC, "libgnc_trace.so")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
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
trace = torch.zeros(296 * 64 * 8, dtype=torch.int64, device="cuda")
clf.lib.gnc_debug_trace.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
clf.lib.gnc_debug_trace(clf._h, ctypes.c_void_p(trace.data_ptr()))
clf.classify(batch, out)
torch.cuda.synchronize()
t = trace.cpu().numpy().reshape(296, 64, 8).astype(np.int64)
valid = t[:, :, 0] > 0
t0 = t[:, :, 0][valid].min()
names = ["full->walk_end", "walk_end->post_done", "post_done->emit_done", "emit_done->iter_end", "iter period", "cnt->lookback_done",
         "walk_end->scan_start"]
it_max = int(valid.sum(1).max())
print("iterations per CTA: min", int(valid.sum(1).min()), "max", it_max, " kernel span (us)",
      (t[:, :, 4].max() - t0) / 1e3)
for it in range(min(it_max, 12)):
    m = valid[:, it] & (t[:, it, 4] > 0)
    if not m.any():
        continue
    a = t[m, it]
    d = [a[:, 1] - a[:, 0], a[:, 2] - a[:, 1], a[:, 3] - a[:, 2], a[:, 4] - a[:, 3]]
    per = (t[m, it + 1, 0] - a[:, 0]) if it + 1 < 64 else a[:, 0] * 0
    per = per[per > 0]
    lb = a[:, 6] - a[:, 5]
    ss = a[:, 5] - a[:, 1]
    row = [f"it{it:2d} start(us) med {np.median(a[:, 0] - t0) / 1e3:7.1f} spread {(a[:, 0].max() - a[:, 0].min()) / 1e3:6.1f} |"]
    for nm, x in zip(names, d + [per, lb, ss]):
        if len(x):
            row.append(f"{nm} med {np.median(x) / 1e3:5.2f} max {x.max() / 1e3:5.2f} |")
    print(" ".join(row))
