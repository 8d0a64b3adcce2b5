"""The native `--json` writer (csrc/fastpack.c: emit_summary_json) against the reference's own text.
Device decisions are stood in for by the C oracle (same output arrays); what is tested is the writer."""
import hashlib

import numpy as np
import pytest

from k8s_gpu_node_checker_b200 import api, engine, native, pnl, synth
from oracle import pnl_oracle as po


@pytest.fixture(scope="module", autouse=True)
def built():
    native.build_fastpack()
    assert pnl._fastpack() is not None


def _result(nodes):
    p = pnl.pack_nodes(nodes)
    o = po.classify(p.rows, p.row_ptr, p.n_nodes)
    defer = np.array([(n, offs, lens) for n, offs, lens in o["defer"]], dtype=engine._DEFER_DT)
    return engine.finish(o["flags"], o["gpu_idx"], o["gpu_total"], o["gpu_vals"], o["counts"], defer, p.rows)


def test_readme_fixture_text(golden):
    from oracle.make_golden import readme_nodes
    nodes = readme_nodes()
    assert api.summary_json_from_result(nodes, _result(nodes)) + "\n" == golden("readme_fixture.json")["json_stdout"]


def test_small_fleets_text(golden):
    g2 = synth.g2_templates()
    for name, rec in golden("fleets.json")["small"].items():
        nodes = [synth.build_node(g2[tid], p) for tid, p in rec["items"]]
        assert api.summary_json_from_result(nodes, _result(nodes)) + "\n" == rec["json_stdout"], name


@pytest.mark.parametrize("name", ["g1_seed0_n10000", "g2_seed0_n10000"])
def test_10k_documents(golden, name):
    rec = golden("fleets.json")["big"][name]
    templates, tid = synth.fleet(rec["kind"], rec["seed"], 0, rec["n"])
    nodes = [synth.build_node(templates[t], synth.g1_payload(i, templates[t]) if t < 1224 else {"name": f"adv-{i}"})
             for i, t in enumerate(tid)]
    txt = api.summary_json_from_result(nodes, _result(nodes)) + "\n"
    assert hashlib.sha256(txt.encode("utf-8", "surrogatepass")).hexdigest() == rec["sha256"]


def test_strings_needing_escapes_and_fallback():
    from kubernetes import client as kc
    import json
    cap = {"nvidia.com/gpu": "2"}
    mk = lambda name, labels, taints: kc.V1Node(   # noqa: E731
        metadata=kc.V1ObjectMeta(name=name, labels=labels), spec=kc.V1NodeSpec(taints=taints),
        status=kc.V1NodeStatus(capacity=cap, conditions=[kc.V1NodeCondition(type="Ready", status="True")]))
    nodes = [mk('q"uo\\te\n\t\x01\x1f é 日本', {"a/b": "x\ry", "ключ": "значение\x7f"}, [kc.V1Taint(key="k", value=None, effect="NoSchedule")]),
             mk(None, {}, []), mk("n3", None, None)]
    res = _result(nodes)
    g, r = api.build_records(nodes, res)
    assert api.summary_json_from_result(nodes, res) == api.summary_json(g, r)
    odd = [mk("n", {"k": 1.5}, None)]                      # a float label: the writer hands over to json.dumps
    res = _result(odd)
    g, r = api.build_records(odd, res)
    assert api.summary_json_from_result(odd, res) == json.dumps({"total_nodes": 1, "ready_nodes": 1, "nodes": g},
                                                                ensure_ascii=False, indent=2)
