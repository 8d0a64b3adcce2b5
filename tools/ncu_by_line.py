"""Aggregate an ncu SASS source page by CUDA source line.

    python tools/ncu_by_line.py <report.ncu-rep> <libgnc.so> [kernel-substring [ncu-kernel-regex]]

ncu's CSV source page is per SASS instruction; nvdisasm -g gives the file:line of each SASS
instruction of the same cubin.  Both list the kernel's instructions in address order, so they are
joined by position.  Prints instructions executed and stall samples per source line.
"""
import csv
import re
import subprocess
import sys
import tempfile
import os
from collections import defaultdict


def sass_lines(so, kernel_sub):
    d = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=d, capture_output=True)
    cubin = [os.path.join(d, f) for f in os.listdir(d) if f.endswith(".cubin")][0]
    txt = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
    out, cur, active = [], ("?", 0), False
    for ln in txt.splitlines():
        m = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
        if m:
            active = kernel_sub in m.group(1)
            continue
        if not active:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", ln):
            out.append(cur)
    return out


def main():
    rep, so = sys.argv[1], sys.argv[2]
    ksub = sys.argv[3] if len(sys.argv) > 3 else "classify_kernelILi224"
    kfilter = ["-k", "regex:" + sys.argv[4]] if len(sys.argv) > 4 else []   # reports holding several kernels: pick one
    csvtxt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"] + kfilter, capture_output=True, text=True).stdout
    rows = list(csv.reader(csvtxt.splitlines()))
    hdr = rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    data = rows[2:]
    lines = sass_lines(so, ksub)
    if len(lines) != len(data):
        print(f"warning: {len(lines)} SASS instructions in the cubin vs {len(data)} in the report", file=sys.stderr)
    agg = defaultdict(lambda: [0, 0, defaultdict(int)])
    stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    for (f, l), r in zip(lines, data):
        a = agg[(f, l)]
        a[0] += int(r[ix["Instructions Executed"]] or 0)
        a[1] += int(r[ix["# Samples"]] or 0)
        for h in stall_cols:
            a[2][h] += int(r[ix[h]] or 0)
    src = {}
    ti = sum(a[0] for a in agg.values())
    ts = sum(a[1] for a in agg.values())
    print(f"total warp-instructions {ti}, samples {ts}")
    print("--- by samples ---")
    for (f, l), a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:25]:
        top = sorted(a[2].items(), key=lambda kv: -kv[1])[:2]
        print(f"{f}:{l:<5} samples {a[1]:>6} ({100*a[1]/ts:4.1f}%) inst {a[0]:>9}  {top}  {line_text(src, f, l)}")
    print("--- by instructions ---")
    for (f, l), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:25]:
        print(f"{f}:{l:<5} inst {a[0]:>9} ({100*a[0]/ti:4.1f}%) samples {a[1]:>6}  {line_text(src, f, l)}")


def line_text(cache, f, l):
    if f not in cache:
        cache[f] = []
        for root in ("k8s-gpu-node-checker_b200/csrc", "include", "."):
            p = os.path.join(root, f)
            if os.path.exists(p):
                cache[f] = open(p).read().splitlines()
                break
    return cache[f][l - 1].strip()[:90] if 0 < l <= len(cache[f]) else ""


if __name__ == "__main__":
    main()
