#!/usr/bin/env python
"""Benchmark of the node-classification path (BASELINE.json metric: node records/sec at 10M nodes).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

One step = one pass of the hot path over the whole synthetic G1 fleet (10M nodes), sharded
contiguously over the ranks: each rank classifies its shard (one kernel launch through the C ABI),
then for N > 1 the ranks combine counts and gather the compacted indices.  Inputs are resident in HBM
before the timed region (2 GB of rows >> the 126 MB L2, so no step sees a warm cache).  After the timed
region every rank checks what it holds -- flags, the fleet-wide ordered index list, the counts -- against
the reference's per-template outputs (tests/golden/templates.json) and rank 0 prints ONE JSON line
(key list at the bottom); a mismatch makes the run fail.

At N = 1 the same process then measures, each as its own object in the line:
  e2e        the same metric through gnc_classify_host (host buffers in and out)
  json       BASELINE.json configs[4]: a 1M-node /api/v1/nodes document -> GPU structural scan -> classify
  e2e_seam   the reference's seam itself: list_gpu_nodes(api) on V1Node objects, ours next to the reference's
  cpu_baseline   the reference's loop on one host core

`--impl reference` times the reference's own `list_gpu_nodes` (check-gpu-node.py:215-226, loaded from the
compiled copy oracle/build_ref.py leaves in oracle/_ref/; the oracle's restatement if that is missing) over
the full 10M-node fleet per step, in 250k-node chunks on all host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FLEET_NODES = 10_000_000
METRIC = "node_records_per_sec"
UNIT = "nodes/s"
WORKLOAD = "g1-10M-node synthetic NodeList (BASELINE.json configs[2]; seed 0)"
JSON_WORKLOAD = "g1-1M-node raw /api/v1/nodes JSON document (BASELINE.json configs[4]; seed 0)"
CACHE_NOTE = "per-step inputs (2.07 GB of rows at 10M nodes) exceed the 126 MB L2; no flush needed between steps"
TRAFFIC_FILE = os.path.join(ROOT, "profiles", "classify_traffic.json")
STUB_DIR = os.path.join(ROOT, "tests", "k8s_stub")


def fleet_config(n_total: int) -> dict:
    """The `config` object: the same in both arms (it names the workload, nothing about how an arm ran it)."""
    return {"workload": WORKLOAD if n_total == FLEET_NODES else f"g1-{n_total}-node synthetic NodeList (seed 0)",
            "nodes": n_total, "seed": 0, "cache": CACHE_NOTE}


# --------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons, sampled while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self._stop, self._t = index, [], threading.Event(), None

    def _run(self):
        try:
            p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                  "-i", str(self.index), "-lms", "100"], stdout=subprocess.PIPE, text=True)
        except OSError:
            return
        self._p = p
        for line in p.stdout:
            self.rows.append([x.strip() for x in line.split(",")])
            if self._stop.is_set():
                break
        p.terminate()

    def start(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        time.sleep(0.35)

    def stop(self):
        time.sleep(0.15)
        self._stop.set()
        if self._t:
            self._t.join(timeout=3)
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 7 for n, v in zip(names, r[3:7]) if v.lower().startswith("active")})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def measured_traffic(variant: str, n_nodes: int):
    """dram__bytes_read.sum + dram__bytes_write.sum of one classify launch, from the committed ncu capture,
    if that capture is of this kernel variant on this workload; else (None, why)."""
    try:
        with open(TRAFFIC_FILE) as f:
            t = json.load(f)
    except Exception as e:
        return None, f"no capture on file ({e.__class__.__name__})"
    if t.get("variant") != variant or int(t.get("nodes", -1)) != n_nodes:
        return None, f"{os.path.relpath(TRAFFIC_FILE, ROOT)} is a capture of {t.get('variant')} at {t.get('nodes')} nodes, this run is {variant} at {n_nodes}"
    return int(t["dram_bytes_read"]) + int(t["dram_bytes_write"]), f"{os.path.relpath(TRAFFIC_FILE, ROOT)} <- {t.get('source')}"


def bind_near_gpu(torch, index: int):
    """Keep this rank's threads (and so the first touch of its pinned buffers) on the CPUs next to its GPU."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByUUID("GPU-" + str(torch.cuda.get_device_properties(index).uuid))
        words = pynvml.nvmlDeviceGetCpuAffinity(h, ((os.cpu_count() or 64) + 63) // 64)
        cpus = [64 * i + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1]
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return None


# --------------------------------------------------------------------------------------------------
# Truth: the reference's own per-template outputs (tests/golden/templates.json, written by oracle/make_golden.py
# from check-gpu-node.py's extract_node_info / list_gpu_nodes), expanded by template id.
# --------------------------------------------------------------------------------------------------
KEYS = ("nvidia.com/gpu", "amd.com/gpu", "gpu.intel.com/i915", "intel.com/gpu")


def g1_truth():
    import numpy as np
    with open(os.path.join(ROOT, "tests", "golden", "templates.json")) as f:
        gold = json.load(f)
    recs = gold["records"][:gold["n_g1"]]
    flags = np.zeros(len(recs), np.uint8)
    vals = np.zeros((len(recs), 4), np.int64)
    for i, r in enumerate(recs):
        f_ = 2 if r["ready"] else 0
        for k, key in enumerate(KEYS):
            if key in r["bd"]:
                f_ |= 8 << k
                vals[i, k] = r["bd"][key]
        if r["gpus"] > 0:
            f_ |= 1
        assert int(vals[i].sum()) == r["gpus"]
        flags[i] = f_
    return flags, vals


class Truth:
    """Expected device outputs for nodes [lo, hi) of the seed-0 G1 fleet, on `dev`."""

    def __init__(self, torch, synth, dev):
        flags, vals = g1_truth()
        self.torch, self.synth, self.dev = torch, synth, dev
        self.flags_t = torch.from_numpy(flags).to(dev)
        self.vals_t = torch.from_numpy(vals).to(dev)

    def expect(self, lo: int, n: int):
        torch = self.torch
        tid = torch.from_numpy(self.synth.g1_tid(0, lo, n)).to(self.dev, torch.int64)
        flags = self.flags_t[tid]
        sel = torch.nonzero(flags & 1).flatten()
        vals = self.vals_t[tid[sel]]
        counts = torch.zeros(16, dtype=torch.int64, device=self.dev)
        counts[0] = sel.numel()
        counts[1] = int(((flags[sel] & 2) != 0).sum())
        counts[3] = n
        counts[4:8] = vals.sum(0)                   # G1 totals are far inside int64: the high words stay 0
        return {"flags": flags, "gpu_idx": (sel + lo).to(torch.int32), "gpu_vals": vals, "gpu_total": vals.sum(1), "counts": counts}


def same(torch, got, want) -> bool:
    return got.shape == want.shape and bool(torch.equal(got, want))


# --------------------------------------------------------------------------------------------------
# CPU legs.  These, and only these, may load the reference (oracle/_ref) or the oracle.
# --------------------------------------------------------------------------------------------------
def cpu_classifier():
    """(kind, fn): fn(nodes) -> (gpu_nodes, ready_gpu_nodes) runs the reference's own list_gpu_nodes behind a stubbed
    CoreV1Api.list_node() (kind "reference"), or the oracle's restatement of it when oracle/_ref is absent ("port")."""
    from oracle import reference_loader
    if reference_loader.reference_available():
        ref = reference_loader.load_reference()
        from kubernetes import client as kc

        def run(nodes):
            kc.CoreV1Api.set_fleet(nodes)
            try:
                return ref.list_gpu_nodes(kc.CoreV1Api())          # check-gpu-node.py:215-226
            finally:
                kc.CoreV1Api.set_fleet(None)
        return "reference", run
    from oracle import cgn_oracle
    return "port", cgn_oracle.classify_fleet


def template_nodes():
    """One stub V1Node per G1 template: the CPU legs index these by template id."""
    from oracle import ensure_kubernetes_importable
    ensure_kubernetes_importable()
    from k8s_gpu_node_checker_b200 import synth
    tmpl = synth.g1_templates()
    return [synth.build_node(t, {"name": f"tmpl-{i}"}) for i, t in enumerate(tmpl)]


def cpu_rate(n_sample: int, seed: int = 0):
    """The reference's loop on the first `n_sample` nodes of the same fleet, one core. nodes/s."""
    from k8s_gpu_node_checker_b200 import synth
    kind, run = cpu_classifier()
    objs = template_nodes()
    tid = synth.g1_tid(seed, 0, n_sample)
    nodes = [objs[t] for t in tid]
    t0 = time.perf_counter()
    g, r = run(nodes)
    dt = time.perf_counter() - t0
    return kind, n_sample / dt, dt, len(g), len(r)


_W = {}


def _ref_worker(args):
    lo, n = args
    from k8s_gpu_node_checker_b200 import synth
    if not _W:
        _W["kind"], _W["run"] = cpu_classifier()
        _W["objs"] = template_nodes()
    tid = synth.g1_tid(0, lo, n)
    nodes = [_W["objs"][t] for t in tid]
    t0 = time.perf_counter()
    g, r = _W["run"](nodes)
    return time.perf_counter() - t0, len(g), len(r), _W["kind"]


def run_reference(args):
    """The reference arm: the reference's classification loop over the whole fleet per step, 250k-node chunks spread
    over all host cores (multiprocessing; the reference itself is single-threaded)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import multiprocessing as mp
    cores = os.cpu_count() or 1
    n_total = args.nodes
    chunk = 250_000                  # small enough that 10M nodes spread evenly over 8..64 processes
    jobs = [(lo, min(chunk, n_total - lo)) for lo in range(0, n_total, chunk)]
    ctx = mp.get_context("fork")
    g = r = 0
    kind = "port"
    with ctx.Pool(min(cores, len(jobs))) as pool:
        for _ in range(args.warmup):
            pool.map(_ref_worker, jobs, chunksize=1)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            res = pool.map(_ref_worker, jobs, chunksize=1)
        dt = time.perf_counter() - t0
        g, r, kind = sum(x[1] for x in res), sum(x[2] for x in res), res[0][3]
    used = min(cores, len(jobs))
    value = n_total * args.steps / dt
    sample = (f"the whole fleet per step: {len(jobs)} chunks of {chunk} nodes over {used} processes "
              f"(host has {cores} cores; python {sys.version.split()[0]})")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int64",
            "data": "synthetic", "config": fleet_config(n_total),
            "detail": {"gpu_nodes": g, "ready_gpu_nodes": r, "processes": used},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": used, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


# --------------------------------------------------------------------------------------------------
# Config 5: raw JSON -> structural scan -> classify (N = 1, after the main leg).
# --------------------------------------------------------------------------------------------------
def json_document(torch, synth, n: int, dev):
    """The seed-0 G1 fleet's first n nodes as one NodeList document, assembled on the device from the 1224
    per-template texts (the payload -- name, labels, taints -- is the template's; the path does not read it)."""
    tmpl = synth.g1_templates()
    texts = [(synth.node_json(t, synth.g1_payload(0, t)) + ",").encode() for t in tmpl]
    tid = synth.g1_tid(0, 0, n)
    body, _, total = synth.expand(texts, torch.from_numpy(tid).to(dev))
    head = b'{"kind":"NodeList","apiVersion":"v1","metadata":{"resourceVersion":"1"},"items":['
    parts = [torch.frombuffer(bytearray(head), dtype=torch.uint8).to(dev), body[:total - 1],
             torch.frombuffer(bytearray(b"]}"), dtype=torch.uint8).to(dev)]
    return torch.cat(parts), tid


def json_cpu_baseline(synth, n_sample: int):
    """json.loads + the client-style adapter + the reference's loop on an n_sample-node document, one core."""
    from oracle import json_adapter
    kind, run = cpu_classifier()
    tmpl = synth.g1_templates()
    tid = synth.g1_tid(0, 0, n_sample)
    cache = {}
    parts = []
    for t in tid:
        t = int(t)
        if t not in cache:
            cache[t] = synth.node_json(tmpl[t], synth.g1_payload(0, tmpl[t]))
        parts.append(cache[t])
    doc = ('{"kind":"NodeList","apiVersion":"v1","metadata":{"resourceVersion":"1"},"items":[' + ",".join(parts) + "]}").encode()
    t0 = time.perf_counter()
    nodes = json_adapter.nodes_from_json(doc)
    t1 = time.perf_counter()
    g, r = run(nodes)
    t2 = time.perf_counter()
    return {"value": n_sample / (t2 - t0), "unit": UNIT, "cores": 1, "kind": kind,
            "sample": f"{n_sample}-node document ({len(doc)} B): json.loads + adapter {t1 - t0:.2f} s, classification loop {t2 - t1:.2f} s",
            "gpu_nodes": len(g), "ready_gpu_nodes": len(r)}


def run_json_leg(args, torch, engine, synth, clf, dev, truth, peak, peak_src):
    import numpy as np
    n = args.json_nodes
    doc, tid = json_document(torch, synth, n, dev)
    nbytes = doc.numel()
    rows = torch.empty(nbytes // 4 + 4096, dtype=torch.uint8, device=dev)     # PNL rows are ~1/6 of the document
    row_ptr = torch.empty(n + 8, dtype=torch.int32, device=dev)
    info = torch.zeros(8, dtype=torch.int64, device=dev)
    out = engine.DeviceResult.alloc(n, dev)

    def scan():
        clf.json_to_pnl_into(doc, n, rows, row_ptr, info)

    scan()
    torch.cuda.synchronize()
    h = info.cpu()
    if int(h[2]):
        raise SystemExit(f"json leg: gnc_json_to_pnl status {int(h[2])}")
    n_nodes, rows_bytes = int(h[0]), int(h[1])
    batch = engine.DeviceBatch(rows, row_ptr, n_nodes, rows_bytes)
    for _ in range(2):
        scan()
        clf.classify(batch, out)
    torch.cuda.synchronize()
    sampler = ClockSampler(dev.index)
    sampler.start()
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.json_steps)]
    l0 = clf.launches
    torch.cuda.synchronize()
    for e in ev:
        e[0].record()
        scan()
        e[1].record()
        clf.classify(batch, out)
        e[2].record()
    torch.cuda.synchronize()
    launches = clf.launches - l0
    clocks = sampler.stop()
    scan_ms = statistics.mean(e[0].elapsed_time(e[1]) for e in ev)
    cls_ms = statistics.mean(e[1].elapsed_time(e[2]) for e in ev)
    # parity of what the scan fed the classifier: node count, and every output against the reference's truth
    exp = truth.expect(0, n)
    g = int(exp["counts"][0])
    ok_counts = n_nodes == n and same(torch, out.counts[:13], exp["counts"][:13]) and int(out.counts[13]) == 0 and int(out.counts[14]) == 0
    ok_idx = same(torch, out.gpu_idx[:g], exp["gpu_idx"]) and same(torch, out.flags[:n], exp["flags"]) and \
        same(torch, out.gpu_vals[:g], exp["gpu_vals"])
    algo = nbytes + rows_bytes + 4 * (n_nodes + 1)                 # document read once, rows and row_ptr written once
    achieved = algo / (scan_ms * 1e-3) / 1e9

    # e2e: the document starts in pinned HOST memory; counts and the ordered index list come back to the host
    e2e = None
    if args.e2e_steps > 0:
        h_doc = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
        h_doc.copy_(doc)
        h_counts = torch.empty(16, dtype=torch.int64).pin_memory()
        h_idx = torch.empty(n, dtype=torch.int32).pin_memory()
        d_doc = torch.empty_like(doc)

        def e2e_step():
            d_doc.copy_(h_doc, non_blocking=True)
            clf.json_to_pnl_into(d_doc, n, rows, row_ptr, info)
            clf.classify(batch, out)
            h_counts.copy_(out.counts, non_blocking=True)
            torch.cuda.current_stream().synchronize()
            gg = int(h_counts[0])
            h_idx[:gg].copy_(out.gpu_idx[:gg], non_blocking=True)
            torch.cuda.current_stream().synchronize()
            return gg
        e2e_step()
        t0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            gg = e2e_step()
        dt = time.perf_counter() - t0
        assert gg == g and np.array_equal(h_idx[:g].numpy(), exp["gpu_idx"].cpu().numpy())
        e2e = {"value": n * args.e2e_steps / dt, "unit": UNIT, "h2d_bytes_per_step": int(nbytes), "d2h_bytes_per_step": int(128 + 4 * g),
               "steps": args.e2e_steps, "ms_per_step": dt / args.e2e_steps * 1e3,
               "api": "pinned host document -> device, gnc_json_to_pnl + gnc_classify, counts and ordered indices back to pinned host memory"}
        del h_doc, d_doc
    cpu = json_cpu_baseline(synth, args.json_cpu_sample) if args.json_cpu_sample > 0 else None
    return {"metric": METRIC, "value": n / ((scan_ms + cls_ms) * 1e-3), "unit": UNIT, "steps": args.json_steps,
            "ms_per_step": scan_ms + cls_ms, "scan_ms": scan_ms, "classify_ms": cls_ms,
            "config": {"workload": JSON_WORKLOAD if n == 1_000_000 else f"g1-{n}-node raw JSON document (seed 0)", "nodes": n,
                       "json_bytes": int(nbytes), "pnl_rows_bytes": rows_bytes,
                       "cache": "the document (1.35 GB at 1M nodes) exceeds the 126 MB L2; no flush needed"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                         "peak_source": peak_src, "kernel": "gnc_json_to_pnl (index, events, measure, write kernels)",
                         "kernel_ms": scan_ms, "algorithmic_bytes_per_launch": int(algo),
                         "note": "algorithmic bytes = document + PNL rows + row_ptr, over the whole scan (all its kernels)"},
            "general_parser_nodes": int(h[5]), "gpu_launches": int(launches), "clocks": clocks,
            "verified": {"counts": bool(ok_counts), "indices": bool(ok_idx)}, "e2e": e2e, "cpu_baseline": cpu}


# --------------------------------------------------------------------------------------------------
# The reference's seam: list_gpu_nodes(api) on V1Node objects (N = 1).
# --------------------------------------------------------------------------------------------------
def run_seam_leg(args, synth, clf):
    if STUB_DIR not in sys.path:
        try:
            import kubernetes  # noqa: F401
        except ImportError:
            sys.path.insert(0, STUB_DIR)
    from kubernetes import client as kc
    from k8s_gpu_node_checker_b200 import api
    n = args.seam_nodes
    tmpl = synth.g1_templates()
    tid = synth.g1_tid(0, 0, n)
    nodes = [synth.build_node(tmpl[t], synth.g1_payload(i, tmpl[t])) for i, t in enumerate(tid)]      # distinct objects, as the client returns
    kc.CoreV1Api.set_fleet(nodes)
    try:
        stub = kc.CoreV1Api()
        api.list_gpu_nodes(stub, clf)                                  # warm-up (staging buffers)
        ours = []
        for _ in range(3):
            t0 = time.perf_counter()
            g, r = api.list_gpu_nodes(stub, clf)
            ours.append(time.perf_counter() - t0)
        t0 = time.perf_counter()
        txt = api.summary_json_from_result(*api.classify_for_summary(nodes, clf))
        t_json = time.perf_counter() - t0
        kind, run = cpu_classifier()
        kc.CoreV1Api.set_fleet(nodes)
        t0 = time.perf_counter()
        g_ref, r_ref = run(nodes)
        t_ref = time.perf_counter() - t0
        t0 = time.perf_counter()
        txt_ref = json.dumps({"total_nodes": len(g_ref), "ready_nodes": len(r_ref), "nodes": g_ref}, ensure_ascii=False, indent=2)
        t_ref_json = time.perf_counter() - t0
    finally:
        kc.CoreV1Api.set_fleet(None)
    same_lists = g == g_ref and r == r_ref and all(a is b for a, b in zip(r, (x for x in g if x["ready"])))
    best = min(ours)
    return {"value": n / best, "unit": UNIT, "nodes": n, "ms": best * 1e3, "api": "api.list_gpu_nodes(api): V1Node objects in, two lists of dicts out",
            "reference": {"value": n / t_ref, "ms": t_ref * 1e3, "kind": kind, "cores": 1},
            "summary_json": {"ms": t_json * 1e3, "reference_ms": (t_ref + t_ref_json) * 1e3, "identical": txt == txt_ref,
                             "api": "classify + native --json writer (summary_json_from_result) vs the reference's loop + json.dumps"},
            "verified": bool(same_lists)}


# --------------------------------------------------------------------------------------------------
def run_b200(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from k8s_gpu_node_checker_b200 import engine, native, synth

    rank = int(os.environ.get("RANK", "0"))
    if not os.path.exists(native.LIB_PATH):      # normally built by __graft_entry__.build(); there is no CPU fallback
        if int(os.environ.get("LOCAL_RANK", "0")) == 0:
            native.build()
        else:
            while not os.path.exists(native.LIB_PATH):
                time.sleep(0.5)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus N > 1 must be launched with torch.distributed.run (one rank per GPU)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    near = bind_near_gpu(torch, local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    n_total = args.nodes
    lo, hi = rank * n_total // world, (rank + 1) * n_total // world      # contiguous shards, fleet order
    n = hi - lo

    # ---- the shard, built in HBM ----
    tmpl = synth.g1_templates()
    rows_t = synth.spec_rows(tmpl)
    tid = synth.g1_tid(0, lo, n)
    rows, row_ptr, total = synth.expand(rows_t, torch.from_numpy(tid).to(dev))
    batch = engine.DeviceBatch(rows, row_ptr, n, total, node_base=lo)

    nccl_id = None
    if world > 1:
        box = [engine.Classifier.comm_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        nccl_id = box[0]
    clf = engine.Classifier(local, rank, world, nccl_id)
    clf.reserve(n)
    gather_kind = "none"
    if world > 1:
        gather_kind = "nccl"
        if args.gather == "fused":
            try:
                clf.fleet_enable(n_total)
                gather_kind = "fused"
            except Exception as e:          # peer mapping unavailable (on every rank alike): the NCCL gather stays in place
                if rank == 0:
                    print(f"fused gather unavailable ({e}); using NCCL", file=sys.stderr)
    out = engine.DeviceResult.alloc(n, dev)
    # the gather target holds the fleet-wide ordered index list and counts (north star: all-reduce of the
    # counts + all-gather-v of the compacted indices); per-node values stay on the rank that owns them
    fleet = engine.DeviceResult.alloc(n_total, dev, defer_cap=1, records=False) if world > 1 else None
    if gather_kind == "fused":
        clf.fleet_target(fleet)      # the gather target is known before the launches: classify + gather are one kernel

    def step():
        clf.classify(batch, out)
        if world > 1:
            clf.gather(out, fleet)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    warm = max(args.warmup, 3)
    for _ in range(warm):
        step()
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    launches0 = clf.launches
    if gather_kind == "fused":
        clf.fleet_timing()          # clears the in-kernel timing sums of the warm-up
    k_ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    t_start, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    if world > 1:
        # A device-side rendezvous between the barrier and the timed region: the ranks' CPUs leave the barrier tens of
        # microseconds apart, and with a step of ~0.1 ms that start-up skew would be charged to the first step of the
        # slower ranks' peers (the fused gather waits for every rank).  Behind this all-reduce every rank's stream
        # resumes at the same moment, with the timed launches already queued.
        dist.all_reduce(torch.zeros(1, device=dev))
    # Per-launch events (the roofline's kernel time) cost a few microseconds of stream time each.  At N = 1 that is
    # under 1 % of a step and they stay; with the fused gather at N > 1 a step is one ~0.1 ms launch whose parts are
    # timed inside the kernel (fused_gather below), so the timed region holds nothing but the launches.
    per_launch = not (world > 1 and gather_kind == "fused")
    t_start.record()
    for i in range(args.steps):
        if per_launch:
            k_ev[i][0].record()
        clf.classify(batch, out)
        if per_launch:
            k_ev[i][1].record()
        if world > 1:
            clf.gather(out, fleet)
    t_end.record()
    barrier()
    total_ms = t_start.elapsed_time(t_end)
    if per_launch:
        kern_ms = [a.elapsed_time(b) for a, b in k_ev]
        # what follows each classify launch on the stream until the next one starts: the gather (N > 1) and launch gaps
        tail_ms = [k_ev[i][1].elapsed_time(k_ev[i + 1][0]) for i in range(args.steps - 1)] or [0.0]
    else:
        kern_ms, tail_ms = [total_ms / args.steps], [0.0]
    launches = clf.launches - launches0
    clocks = sampler.stop() if sampler else None
    fleet_timing = None
    if gather_kind == "fused":
        ft = clf.fleet_timing()
        t = torch.tensor([ft["wait_us"], ft["rebuild_us"], ft["rebuild_us_max"], ft["classify_us"]], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        fleet_timing = {"rebuild_header_us": ft["rebuild_header_us"], "rebuild_groups_us_warp0": ft["rebuild_groups_us"], "classify_us": float(t[3]), "flag_wait_us": float(t[0]), "rebuild_us": float(t[1]), "rebuild_us_longest": float(t[2]),
                        "note": "in-kernel timestamps over the timed steps, slowest rank: what a CTA spends on average classifying its tiles, "
                                "then waiting for every rank's flag, then rebuilding the fleet-wide list"}
    if world > 1:
        t = torch.tensor([total_ms, statistics.mean(kern_ms), statistics.mean(tail_ms)], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms, kern_mean, tail_mean = float(t[0]), float(t[1]), float(t[2])
    else:
        kern_mean, tail_mean = statistics.mean(kern_ms), statistics.mean(tail_ms)

    # ---- parity of what was just timed (outside the timed region), on every rank ----
    truth = Truth(torch, synth, dev)
    mine = truth.expect(lo, n)
    g_local = int(mine["counts"][0])
    ok_local = same(torch, out.flags[:n], mine["flags"]) and same(torch, out.gpu_idx[:g_local], mine["gpu_idx"]) and \
        same(torch, out.gpu_vals[:g_local], mine["gpu_vals"]) and same(torch, out.gpu_total[:g_local], mine["gpu_total"]) and \
        same(torch, out.counts[:13], mine["counts"][:13])
    if world > 1:
        whole = truth.expect(0, n_total)
        g_fleet = int(whole["counts"][0])
        ok_idx = same(torch, fleet.gpu_idx[:g_fleet], whole["gpu_idx"])
        ok_counts = same(torch, fleet.counts[:13], whole["counts"][:13])
        tail = fleet.counts[13:15]
        del whole
    else:
        g_fleet = g_local
        ok_idx = same(torch, out.gpu_idx[:g_local], mine["gpu_idx"])
        ok_counts = same(torch, out.counts[:13], mine["counts"][:13])
        tail = out.counts[13:15]
    ok_clean = int(tail[0]) == 0 and int(tail[1]) == 0 and int(out.counts[13]) == 0 and int(out.counts[14]) == 0   # a timed-out fused wait poisons these
    oks = torch.tensor([ok_local, ok_idx, ok_counts, ok_clean], device=dev, dtype=torch.int32)
    if world > 1:
        dist.all_reduce(oks, op=dist.ReduceOp.MIN)
    ok_local, ok_idx, ok_counts, ok_clean = (bool(x) for x in oks.tolist())
    verified = {"indices": ok_idx, "counts": ok_counts and ok_clean, "per_rank_outputs": ok_local, "ranks": world, "gather": gather_kind,
                "against": "tests/golden/templates.json (the reference's per-template outputs) expanded by template id, every element"}
    ready_fleet = int((fleet.counts if world > 1 else out.counts)[1])
    del mine
    # algorithmic bytes of one launch on this rank (DESIGN.md "Algorithmic bytes"):
    #   in: row_ptr 4(n+1) + rows T      out: flags n + 44 G + counts 128
    algo_bytes = 4 * (n + 1) + total + n + 44 * g_local + 128
    peak, peak_src = measured_peak_gbs()
    achieved = algo_bytes / (kern_mean * 1e-3) / 1e9
    variant = os.environ.get("GNC_VARIANT") or engine.variant_for(total, n)

    # ---- e2e: host buffers in, host buffers out, through gnc_classify_host ----
    e2e = None
    if args.e2e_steps > 0:
        from k8s_gpu_node_checker_b200 import pnl
        h_rows = torch.empty(rows.numel(), dtype=torch.uint8).pin_memory()
        h_rp = torch.empty(row_ptr.numel(), dtype=torch.int32).pin_memory()
        h_rows.copy_(rows)
        h_rp.copy_(row_ptr)
        torch.cuda.synchronize()
        packed = pnl.PackedNodeList(h_rows.numpy(), h_rp.numpy().view(np.uint32), n)
        hout = engine.host_buffers(n, pinned=True)
        clf.classify_host(packed, node_base=lo, out=hout)           # warm-up: allocates the staging buffers
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            clf.classify_host(packed, node_base=lo, out=hout)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t[0])
        want_idx = truth.expect(lo, n)["gpu_idx"].cpu().numpy().view(np.uint32)
        assert int(hout["counts"][0]) == g_local and np.array_equal(hout["gpu_idx"][:g_local], want_idx), "e2e outputs differ from the truth"
        e2e = {"value": n_total * args.e2e_steps / dt, "unit": UNIT, "h2d_bytes_per_step": int(total + 4 * (n + 1)),
               "d2h_bytes_per_step": int(n + 44 * g_local + 128), "steps": args.e2e_steps,
               "ms_per_step": dt / args.e2e_steps * 1e3,
               "api": "gnc_classify_host (pinned host buffers in and out, chunked H2D/D2H inside the timed region)" +
                      ("; per-rank shards, no cross-rank combine in this leg" if world > 1 else ""),
               "cpus_near_gpu": near}
        del h_rows, h_rp, hout, packed

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0 if all((ok_local, ok_idx, ok_counts, ok_clean)) else 1

    traffic, traffic_src = measured_traffic(variant, n) if world == 1 else (None, "single-GPU capture only")
    line = {
        "metric": METRIC, "value": n_total * args.steps / (total_ms * 1e-3), "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": warm, "ms_per_step": total_ms / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int64", "data": "synthetic",
        "config": fleet_config(n_total),
        "detail": {"nodes_per_rank": n, "rows_bytes_per_rank": int(total), "gpu_nodes": g_fleet, "ready_gpu_nodes": ready_fleet,
                   "parallelism": f"shard{world}", "kernel_variant": variant,
                   "collectives": {"none": "none",
                                   "nccl": "one ncclAllGather of per-rank (counts block + compacted indices) slots, then an unpack kernel",
                                   "fused": "none, and no second launch: classify stores per-tile GPU-node ballots, prefixes and the counts into every "
                                            "rank's buffer over NVLink (peer stores) while it runs, then the same kernel waits for the ranks' flags and "
                                            "rebuilds the fleet-wide index list locally"}[gather_kind]},
        "verified": verified,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "kernel": f"gnc::classify_kernel ({variant})",
                     "kernel_ms": kern_mean, "after_kernel_ms": tail_mean, "fused_gather": fleet_timing, "algorithmic_bytes_per_launch": int(algo_bytes),
                     "bytes_per_node": algo_bytes / n, "frac_of_8TBps": achieved / 8000.0},
        "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
    }
    cpu = None
    if world == 1:
        if args.json_nodes > 0:
            del rows, row_ptr, batch, out
            torch.cuda.empty_cache()
            line["json"] = run_json_leg(args, torch, engine, synth, clf, dev, truth, peak, peak_src)
        if args.seam_nodes > 0:
            line["e2e_seam"] = run_seam_leg(args, synth, clf)
    if args.cpu_sample > 0 and world == 1:          # the CPU baseline is taken on rank 0 at N = 1 only
        kind, rate, dt, g, r = cpu_rate(args.cpu_sample)
        cpu = {"value": rate, "unit": UNIT, "cores": 1, "kind": kind,
               "sample": f"first {args.cpu_sample} nodes of the same fleet, {dt:.1f} s, python {sys.version.split()[0]}, "
                         f"host has {os.cpu_count()} cores"}
    line["cpu_baseline"] = cpu
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    bad = not all((ok_local, ok_idx, ok_counts, ok_clean))
    for leg in ("json", "e2e_seam"):
        v = line.get(leg, {}).get("verified")
        if v is False or (isinstance(v, dict) and not all(v.values())):
            bad = True
    if bad:
        print("bench: outputs differ from the reference's truth (see the \"verified\" keys)", file=sys.stderr)
    return 1 if bad else 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--nodes", type=int, default=FLEET_NODES)
    ap.add_argument("--gather", default="fused", choices=["fused", "nccl"], help="multi-GPU combine step")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--cpu-sample", type=int, default=2_000_000, help="nodes the 1-core CPU baseline classifies (0 = skip)")
    ap.add_argument("--json-nodes", type=int, default=1_000_000, help="nodes of the raw-JSON leg's document (N = 1 only; 0 = skip)")
    ap.add_argument("--json-steps", type=int, default=5)
    ap.add_argument("--json-cpu-sample", type=int, default=50_000, help="nodes of the document the JSON CPU baseline parses (0 = skip)")
    ap.add_argument("--seam-nodes", type=int, default=200_000, help="V1Node objects of the list_gpu_nodes(api) leg (N = 1 only; 0 = skip)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
