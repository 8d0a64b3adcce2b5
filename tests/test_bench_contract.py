"""bench.py keeps its contract: one JSON line with the agreed keys, on both arms."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
             "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e", "gpu_launches"}


def _run(args, timeout):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, capture_output=True, text=True, timeout=timeout,
                         cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1, out.stdout[-2000:]
    return json.loads(lines[0])


def test_reference_arm_line():
    d = _run(["--impl", "reference", "--steps", "1", "--warmup", "1", "--nodes", "500000"], 600)
    assert BASE_KEYS <= set(d) and d["impl"] == "reference" and d["gpu_launches"] == 0
    from oracle.reference_loader import reference_available
    assert d["cpu_baseline"]["kind"] == ("reference" if reference_available() else "port")
    assert d["cpu_baseline"]["cores"] >= 1 and d["value"] > 1e4 and d["config"]["nodes"] == 500000
    assert d["detail"]["gpu_nodes"] > 0.6 * 500000
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["value"] == d["value"]


@pytest.mark.gpu
def test_b200_arm_line():
    d = _run(["--nodes", "400000", "--steps", "4", "--warmup", "3", "--e2e-steps", "1", "--cpu-sample", "20000",
              "--json-nodes", "30000", "--json-steps", "2", "--json-cpu-sample", "2000", "--seam-nodes", "5000"], 900)
    assert BASE_KEYS | {"roofline", "clocks", "verified", "json", "e2e_seam"} <= set(d)
    assert d["verified"]["indices"] and d["verified"]["counts"] and d["verified"]["per_rank_outputs"] and d["verified"]["ranks"] == 1
    j = d["json"]
    assert j["verified"] == {"counts": True, "indices": True} and j["general_parser_nodes"] == 0 and j["gpu_launches"] > 2
    assert 0 < j["roofline"]["frac"] < 1.5 and j["e2e"]["h2d_bytes_per_step"] == j["config"]["json_bytes"] and j["cpu_baseline"]["value"] > 0
    s = d["e2e_seam"]
    assert s["verified"] and s["summary_json"]["identical"] and s["value"] > 0 and s["reference"]["value"] > 0
    assert d["metric"] == "node_records_per_sec" and d["gpu_launches"] == 4 and d["n_gpus"] == 1
    r = d["roofline"]
    assert r["bound"] == "hbm" and 0 < r["frac"] < 1.5 and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    assert d["e2e"]["h2d_bytes_per_step"] > 0 and d["e2e"]["d2h_bytes_per_step"] > 0 and d["e2e"]["value"] < d["value"]
    assert d["cpu_baseline"]["cores"] == 1 and d["detail"]["gpu_nodes"] > 0.6 * 400000
