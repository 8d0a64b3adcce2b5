"""Generate tests/golden/*.json by running the UNMODIFIED reference (this container only).

    python -m oracle.make_golden

Imports /root/reference/check-gpu-node.py through oracle/reference_loader.py (stub `kubernetes`,
see tests/k8s_stub) and records what ITS functions and ITS main() produce:

  readme_fixture.json   README.md:93-121's 2-node fleet: `--json` text, table text, exit code
  quantity_kat.json     gpu_capacity() (check-gpu-node.py:181-196) on every QUANTITY_CASES value
  templates.json        extract_node_info() (:199-212) + GPU/Ready membership from
                        list_gpu_nodes() (:215-226) for every G1 and G2 template (by spec digest)
  fleets.json           main() `--json` output (:273-279) and exit code (:289-293) for small fleets
                        in full, and digests + counts for 10k-node fleets (configs 2 and 4)
The vectors travel to the GPU box; the reference does not.
"""
from __future__ import annotations

import contextlib
import hashlib
import io
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle.reference_loader import load_reference  # noqa: E402
from k8s_gpu_node_checker_b200 import synth  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def spec_digest(spec) -> str:
    return hashlib.sha256(json.dumps(spec, sort_keys=True, ensure_ascii=True).encode()).hexdigest()[:16]


def run_main(ref, nodes, argv):
    """(stdout text, exit code) of the reference's main() over a stubbed list_node()."""
    from kubernetes import client as kc
    kc.CoreV1Api.set_fleet(nodes)
    out = io.StringIO()
    old = sys.argv
    sys.argv = ["check-gpu-node.py"] + argv
    try:
        with contextlib.redirect_stdout(out):
            code = ref.main()
    finally:
        sys.argv = old
        kc.CoreV1Api.set_fleet(None)
    return out.getvalue(), code


def readme_nodes():
    """README.md:93-97 / :103-121: two Ready nvidia nodes, 4 and 8 GPUs, one label, no taints."""
    mk = lambda name, n: synth.build_node(   # noqa: E731
        {"capacity": [["cpu", "4"], ["memory", "16Gi"], ["nvidia.com/gpu", str(n)], ["pods", "110"]],
         "conditions": [["MemoryPressure", "False"], ["Ready", "True"]]},
        {"name": name, "labels": {"node.kubernetes.io/instance-type": "g4dn.xlarge"}, "taints": None})
    return [mk("gpu-node-1", 4), mk("gpu-node-2", 8)]


def small_fleets():
    """name -> (templates, [(template id, payload)])."""
    g2 = synth.g2_templates()
    adv0 = 1224
    gpu_ready = next(i for i, t in enumerate(g2) if ["nvidia.com/gpu", "8"] in t["capacity"]
                     and t["conditions"][-1] == ["Ready", "True"])
    gpu_notready = next(i for i, t in enumerate(g2) if ["nvidia.com/gpu", "8"] in t["capacity"]
                        and t["conditions"][-1] == ["Ready", "False"])
    no_gpu = 0
    pay = lambda i: {"name": f"n{i}", "labels": {"idx": str(i)} if i % 2 else None,   # noqa: E731
                     "taints": [{"key": "k", "value": None, "effect": "NoSchedule"}] if i % 3 == 0 else None}
    fleets = {
        "empty": [],
        "zero_gpu": [(no_gpu, pay(i)) for i in range(5)],
        "zero_ready": [(gpu_notready, pay(0)), (no_gpu, pay(1)), (gpu_notready, pay(2))],
        "mixed": [(gpu_ready, pay(0)), (no_gpu, pay(1)), (gpu_notready, pay(2)), (gpu_ready, pay(3))],
        "every_adversarial": [(adv0 + j, pay(j)) for j in range(len(g2) - adv0)],
    }
    return g2, fleets


def main():
    os.makedirs(GOLD, exist_ok=True)
    ref = load_reference()
    from kubernetes import client as kc

    # 1. README fixture --------------------------------------------------------------------
    nodes = readme_nodes()
    jtxt, jcode = run_main(ref, nodes, ["--json"])
    ttxt, tcode = run_main(ref, nodes, [])
    with open(os.path.join(GOLD, "readme_fixture.json"), "w") as f:
        json.dump({"json_stdout": jtxt, "json_exit": jcode, "table_stdout": ttxt, "table_exit": tcode},
                  f, ensure_ascii=True, indent=1)

    # 2. quantity known answers ------------------------------------------------------------
    kat = []
    for q in synth.QUANTITY_CASES:
        node = synth.build_node({"capacity": [["nvidia.com/gpu", q]], "conditions": None})
        caps = ref.gpu_capacity(node)
        kat.append({"value": q if not isinstance(q, float) else {"float": q},
                    "type": type(q).__name__, "kept": "nvidia.com/gpu" in caps,
                    "int": caps.get("nvidia.com/gpu")})
    with open(os.path.join(GOLD, "quantity_kat.json"), "w") as f:
        json.dump(kat, f, ensure_ascii=True, indent=0)

    # 3. every template --------------------------------------------------------------------
    tmpl = synth.g2_templates()
    recs = []
    for t in tmpl:
        node = synth.build_node(t)
        info = ref.extract_node_info(node)
        kc.CoreV1Api.set_fleet([node])
        g, r = ref.list_gpu_nodes(kc.CoreV1Api())
        recs.append({"d": spec_digest(t), "ready": info["ready"], "gpus": info["gpus"],
                     "bd": info["gpu_breakdown"], "in_gpu": len(g), "in_ready": len(r)})
    kc.CoreV1Api.set_fleet(None)
    with open(os.path.join(GOLD, "templates.json"), "w") as f:
        json.dump({"n_g1": 1224, "n": len(tmpl), "records": recs}, f, ensure_ascii=True, separators=(",", ":"))

    # 4. fleets ----------------------------------------------------------------------------
    g2, fleets = small_fleets()
    out = {}
    for name, items in fleets.items():
        nodes = [synth.build_node(g2[tid], p) for tid, p in items]
        txt, code = run_main(ref, nodes, ["--json"])
        out[name] = {"items": [[tid, p] for tid, p in items], "json_stdout": txt, "exit": code}
    big = {}
    for kind, seed, n in (("g1", 0, 10000), ("g2", 0, 10000), ("g1", 7, 3000)):
        templates, tid = synth.fleet(kind, seed, 0, n)
        nodes = [synth.build_node(templates[t], synth.g1_payload(i, templates[t]) if t < 1224 else {"name": f"adv-{i}"})
                 for i, t in enumerate(tid)]
        txt, code = run_main(ref, nodes, ["--json"])
        doc = json.loads(txt)
        big[f"{kind}_seed{seed}_n{n}"] = {
            "kind": kind, "seed": seed, "n": n, "exit": code,
            "sha256": hashlib.sha256(txt.encode()).hexdigest(), "chars": len(txt),
            "total_nodes": doc["total_nodes"], "ready_nodes": doc["ready_nodes"],
            "first_names": [d["name"] for d in doc["nodes"][:5]],
            "gpus_sum_hex": hex(sum(d["gpus"] for d in doc["nodes"]))}
    with open(os.path.join(GOLD, "fleets.json"), "w") as f:
        json.dump({"small": out, "big": big}, f, ensure_ascii=True, indent=0)
    print("golden vectors written to", GOLD)


if __name__ == "__main__":
    main()
