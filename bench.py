#!/usr/bin/env python
"""Benchmark of the node-classification path (BASELINE.json metric: node records/sec at 10M nodes).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

One step = one pass of the hot path over the whole synthetic G1 fleet (10M nodes), sharded
contiguously over the ranks: each rank classifies its shard (one kernel launch through the C ABI),
then for N > 1 the ranks combine counts and gather the compacted indices over NCCL.  Inputs are
resident in HBM before the timed region (2 GB of rows >> the 126 MB L2, so no step sees a warm
cache).  Rank 0 prints ONE JSON line (see the key list at the bottom).

`--impl reference` times the reference's own CPU algorithm (oracle/cgn_oracle.py, the restatement
of check-gpu-node.py:215-226 pinned to the reference by tests/golden) on all host cores.
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
# dram__bytes_read.sum + dram__bytes_write.sum of ONE classify launch on the default workload (N=1), from the
# `ncu --set full` capture summarised in profiles/r01_ncu_classify_10m.txt (2.1139 GB read + 0.2978 GB written)
NCU_TRAFFIC_BYTES_DEFAULT_WORKLOAD = 2_411_907_392
METRIC = "node_records_per_sec"
UNIT = "nodes/s"
WORKLOAD = "g1-10M-node synthetic NodeList (BASELINE.json configs[2]; seed 0)"


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


def template_nodes():
    """One stub V1Node per G1 template: the CPU legs index these by template id."""
    from oracle import ensure_kubernetes_importable
    ensure_kubernetes_importable()
    from k8s_gpu_node_checker_b200 import synth
    tmpl = synth.g1_templates()
    return [synth.build_node(t, {"name": f"tmpl-{i}"}) for i, t in enumerate(tmpl)]


def cpu_port_rate(n_sample: int, seed: int = 0):
    """The reference's loop (restated) on `n_sample` nodes of the same fleet, one core. nodes/s."""
    from oracle import cgn_oracle
    from k8s_gpu_node_checker_b200 import synth
    objs = template_nodes()
    tid = synth.g1_tid(seed, 0, n_sample)
    nodes = [objs[t] for t in tid]
    t0 = time.perf_counter()
    g, r = cgn_oracle.classify_fleet(nodes)
    dt = time.perf_counter() - t0
    return n_sample / dt, dt, len(g), len(r)


# --------------------------------------------------------------------------------------------------
def _ref_worker(args):
    lo, n = args
    from oracle import cgn_oracle
    from k8s_gpu_node_checker_b200 import synth
    global _W_OBJS
    try:
        _W_OBJS
    except NameError:
        _W_OBJS = template_nodes()
    tid = synth.g1_tid(0, lo, n)
    nodes = [_W_OBJS[t] for t in tid]
    t0 = time.perf_counter()
    g, r = cgn_oracle.classify_fleet(nodes)
    return time.perf_counter() - t0, len(g), len(r)


def run_reference(args):
    """The reference arm: Python classification loop over all host cores (multiprocessing)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import multiprocessing as mp
    cores = os.cpu_count() or 1
    per_worker = 150_000
    ctx = mp.get_context("fork")
    with ctx.Pool(cores) as pool:
        jobs = [(w * per_worker, per_worker) for w in range(cores)]
        for _ in range(args.warmup):
            pool.map(_ref_worker, jobs)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            pool.map(_ref_worker, jobs)
        dt = time.perf_counter() - t0
    nodes_per_step = per_worker * cores
    value = nodes_per_step * args.steps / dt
    sample = f"{nodes_per_step} nodes/step of the same fleet ({per_worker} per process x {cores} processes)"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int64",
            "data": "synthetic", "config": {"workload": WORKLOAD, "sample": sample},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


# --------------------------------------------------------------------------------------------------
def run_b200(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from oracle import ensure_kubernetes_importable
    ensure_kubernetes_importable()
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
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    n_total = args.nodes
    lo, hi = rank * n_total // world, (rank + 1) * n_total // world      # contiguous shards, fleet order
    n = hi - lo

    # ---- the shard, built in HBM ----
    tmpl = synth.g1_templates()
    rows_t = synth.template_rows(tmpl)
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
            except Exception as e:          # peer mapping unavailable: the NCCL gather stays in place
                if rank == 0:
                    print(f"fused gather unavailable ({e}); using NCCL", file=sys.stderr)
    out = engine.DeviceResult.alloc(n, dev)
    # the gather target holds the fleet-wide ordered index list and counts (north star: all-reduce of the
    # counts + all-gather-v of the compacted indices); per-node values stay on the rank that owns them
    fleet = engine.DeviceResult.alloc(n_total, dev, defer_cap=1, records=False) if world > 1 else None

    def step():
        clf.classify(batch, out)
        if world > 1:
            clf.gather(out, fleet)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    launches0 = clf.launches
    k_ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    t_start, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_start.record()
    for i in range(args.steps):
        k_ev[i][0].record()
        clf.classify(batch, out)
        k_ev[i][1].record()
        if world > 1:
            clf.gather(out, fleet)
    t_end.record()
    barrier()
    total_ms = t_start.elapsed_time(t_end)
    kern_ms = [a.elapsed_time(b) for a, b in k_ev]
    # what follows each classify launch on the stream until the next one starts: the gather (N > 1) and launch gaps
    tail_ms = [k_ev[i][1].elapsed_time(k_ev[i + 1][0]) for i in range(args.steps - 1)] or [0.0]
    launches = clf.launches - launches0
    clocks = sampler.stop() if sampler else None
    if world > 1:
        t = torch.tensor([total_ms, statistics.mean(kern_ms), statistics.mean(tail_ms)], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms, kern_mean, tail_mean = float(t[0]), float(t[1]), float(t[2])
    else:
        kern_mean, tail_mean = statistics.mean(kern_ms), statistics.mean(tail_ms)

    counts = (fleet.counts if world > 1 else out.counts).cpu().numpy()
    g_fleet = int(counts[0])
    local_counts = out.counts.cpu().numpy()
    g_local = int(local_counts[0])
    # algorithmic bytes of one launch on this rank (DESIGN.md "Algorithmic bytes"):
    #   in: row_ptr 4(n+1) + rows T      out: flags n + 44 G + counts 128
    algo_bytes = 4 * (n + 1) + total + n + 44 * g_local + 128
    peak, peak_src = measured_peak_gbs()
    achieved = algo_bytes / (kern_mean * 1e-3) / 1e9

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
        assert int(hout["counts"][0]) == g_local
        e2e = {"value": n_total * args.e2e_steps / dt, "unit": UNIT, "h2d_bytes_per_step": int(total + 4 * (n + 1)),
               "d2h_bytes_per_step": int(n + 44 * g_local + 128), "steps": args.e2e_steps,
               "ms_per_step": dt / args.e2e_steps * 1e3,
               "api": "gnc_classify_host (pinned host buffers in and out, chunked H2D/D2H inside the timed region)"}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    cpu = None
    if args.cpu_sample > 0:
        rate, dt, g, r = cpu_port_rate(args.cpu_sample)
        cpu = {"value": rate, "unit": UNIT, "cores": 1, "kind": "port",
               "sample": f"first {args.cpu_sample} nodes of the same fleet, {dt:.1f} s, python {sys.version.split()[0]}, "
                         f"host has {os.cpu_count()} cores"}
    line = {
        "metric": METRIC, "value": n_total * args.steps / (total_ms * 1e-3), "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": total_ms / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "nodes": n_total, "nodes_per_rank": n, "rows_bytes_per_rank": int(total),
                   "gpu_nodes": g_fleet, "ready_gpu_nodes": int(counts[1]), "parallelism": f"shard{world}",
                   "cache": "inputs (rows_bytes_per_rank) exceed the 126 MB L2; no flush needed",
                   "collectives": {"none": "none",
                                   "nccl": "one ncclAllGather of per-rank (counts block + compacted indices) slots, then an unpack kernel",
                                   "fused": "none after the kernel: classify stores per-tile GPU-node ballots, prefixes and the counts into every "
                                            "rank's buffer over NVLink (peer stores), gather = flag wait + local rebuild of the index list"}[gather_kind]},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": NCU_TRAFFIC_BYTES_DEFAULT_WORKLOAD if (world == 1 and n_total == FLEET_NODES) else None,
                     "traffic_source": "profiles/r01_ncu_classify_10m.txt (ncu --set full, one launch)", "peak_source": peak_src, "kernel": "gnc::classify_kernel",
                     "kernel_ms": kern_mean, "after_kernel_ms": tail_mean, "algorithmic_bytes_per_launch": int(algo_bytes),
                     "bytes_per_node": algo_bytes / n, "frac_of_8TBps": achieved / 8000.0},
        "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
    }
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


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
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
