"""Instruction and stall-sample totals per function/region of classify.cuh from an ncu report.
    python tools/ncu_by_region.py <report.ncu-rep> <libgnc.so> [kernel-substring]"""
import csv
import subprocess
import sys
from collections import defaultdict

sys.path.insert(0, "tools")
import ncu_by_line as m  # noqa: E402

rep, so = sys.argv[1], sys.argv[2]
ksub = sys.argv[3] if len(sys.argv) > 3 else "classify_kernelILi224"
csvtxt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(csvtxt.splitlines()))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
data = rows[2:]
lines = m.sass_lines(so, ksub)
src = open("k8s-gpu-node-checker_b200/csrc/classify.cuh").read().splitlines()
marks = [(i, l.strip()[:72]) for i, l in enumerate(src, 1)
         if l.startswith("__device__") or l.startswith("__global__") or "auto emit_stashed" in l or "producer ---" in l
         or "scan warps:" in l or "// ------------------------------ walkers" in l or "flags, ranks, partial counts" in l
         or "grid-wide finish" in l]


def region(f, l):
    if f != "classify.cuh":
        return f
    r = "?"
    for i, name in marks:
        if i <= l:
            r = name
    return r


agg = defaultdict(lambda: [0, 0])
for (f, l), r in zip(lines, data):
    k = region(f, l)
    agg[k][0] += int(r[ix["Instructions Executed"]] or 0)
    agg[k][1] += int(r[ix["# Samples"]] or 0)
tot = sum(v[0] for v in agg.values())
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:24]:
    print(f"{v[0]:>10} {100 * v[0] / tot:5.1f}%  samples {v[1]:>5}  {k}")
print("total", tot)
