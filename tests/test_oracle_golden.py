"""The Python restatement (oracle/cgn_oracle.py) held to vectors produced by the reference itself."""
import hashlib
import json
import sys

import pytest
from hypothesis import given, settings, strategies as st

from oracle import cgn_oracle as o
from k8s_gpu_node_checker_b200 import synth
from tests.helpers import spec_digest


def _fleet_json_and_exit(nodes):
    g, r = o.classify_fleet(nodes)
    return o.summary_json(g, r) + "\n", o.exit_code(g, r)     # print() adds the newline (cgn.py:279)


def test_readme_fixture(golden):
    """Config 1: README.md:93-121 through the reference's main() == the restatement."""
    from oracle.make_golden import readme_nodes
    gold = golden("readme_fixture.json")
    txt, code = _fleet_json_and_exit(readme_nodes())
    assert txt == gold["json_stdout"]
    assert code == gold["json_exit"] == 0
    doc = json.loads(txt)
    assert [(n["name"], n["gpus"]) for n in doc["nodes"]] == [("gpu-node-1", 4), ("gpu-node-2", 8)]
    assert list(doc["nodes"][0]) == ["name", "ready", "gpus", "gpu_breakdown", "labels", "taints"]


def test_quantity_known_answers(golden):
    kat = golden("quantity_kat.json")
    assert len(kat) == len(synth.QUANTITY_CASES)
    for rec, q in zip(kat, synth.QUANTITY_CASES):
        node = synth.build_node({"capacity": [["nvidia.com/gpu", q]], "conditions": None})
        bd = o.gpu_breakdown_of(node)
        assert ("nvidia.com/gpu" in bd) == rec["kept"], repr(q)[:40]
        assert bd.get("nvidia.com/gpu") == rec["int"], repr(q)[:40]
        if isinstance(q, str):          # the spelled-out grammar agrees too ('' is skipped before int())
            want = rec["int"] if rec["kept"] else None
            assert o.int_grammar(q) == want, repr(q)[:40]


def test_every_template(golden):
    gold = golden("templates.json")
    tmpl = synth.g2_templates()
    assert gold["n"] == len(tmpl) and gold["n_g1"] == 1224 == len(synth.g1_templates())
    for t, rec in zip(tmpl, gold["records"]):
        assert spec_digest(t) == rec["d"], "template definitions drifted; rerun oracle/make_golden.py"
        info = o.node_record(synth.build_node(t))
        assert (info["ready"], info["gpus"], info["gpu_breakdown"]) == (rec["ready"], rec["gpus"], rec["bd"])
        g, r = o.classify_fleet([synth.build_node(t)])
        assert (len(g), len(r)) == (rec["in_gpu"], rec["in_ready"])
        assert list(info["gpu_breakdown"]) == [k for k in o.GPU_KEYS if k in rec["bd"]]    # key order


def test_small_fleets_json_and_exit_codes(golden):
    gold = golden("fleets.json")["small"]
    g2 = synth.g2_templates()
    assert {k: v["exit"] for k, v in gold.items()} == {
        "empty": 2, "zero_gpu": 2, "zero_ready": 3, "mixed": 0, "every_adversarial": 0}
    for name, rec in gold.items():
        nodes = [synth.build_node(g2[tid], p) for tid, p in rec["items"]]
        txt, code = _fleet_json_and_exit(nodes)
        assert txt == rec["json_stdout"], name
        assert code == rec["exit"], name


@pytest.mark.parametrize("name", ["g1_seed0_n10000", "g2_seed0_n10000", "g1_seed7_n3000"])
def test_10k_fleets_digest(golden, name):
    """Configs 2 and 4 at 10k nodes: byte-identical `--json` text (by SHA-256) and exit code."""
    rec = golden("fleets.json")["big"][name]
    templates, tid = synth.fleet(rec["kind"], rec["seed"], 0, rec["n"])
    nodes = [synth.build_node(templates[t], synth.g1_payload(i, templates[t]) if t < 1224 else {"name": f"adv-{i}"})
             for i, t in enumerate(tid)]
    txt, code = _fleet_json_and_exit(nodes)
    assert hashlib.sha256(txt.encode()).hexdigest() == rec["sha256"]
    assert code == rec["exit"]
    g, r = o.classify_fleet(nodes)
    assert (len(g), len(r)) == (rec["total_nodes"], rec["ready_nodes"])
    assert hex(sum(n["gpus"] for n in g)) == rec["gpus_sum_hex"]


# ---- int(str) grammar: the interpreter is the authority ------------------------------------
def _py_int(s):
    try:
        return int(s)
    except ValueError:
        return None


def test_grammar_every_code_point():
    """One code point around a digit: classifies all of Unicode as space / digit / neither."""
    for cp in list(range(0, 0x3100)) + list(range(0xA000, 0xAC00)) + list(range(0xFE00, 0x12000)) + \
            list(range(0x1D700, 0x1FC00)):
        ch = chr(cp)
        for s in (ch, ch + "4", "4" + ch, "1" + ch + "2"):
            assert o.int_grammar(s) == _py_int(s), hex(cp)


def test_unicode_tables_cover_the_database():
    runs, spaces = o.unicode_decimal_runs(), o.unicode_wide_spaces()
    digit_cps = {r + d for r in runs for d in range(10)}
    for cp in range(127, 0x110000):
        v = _py_int(chr(cp))
        assert (v is not None) == (cp in digit_cps), hex(cp)
        assert (cp not in digit_cps and _py_int(chr(cp) + "7") == 7) == (cp in spaces), hex(cp)


_alphabet = st.sampled_from(list("0123456789_+- \t\n\x0b\x0c\r\x1c\x00kKi.e") +
                            ["٤", "４", "\xa0", " ", "\x85", "\U0001d7dc", "²", "\ud800", "​"])


@settings(max_examples=3000, deadline=None)
@given(st.lists(_alphabet, max_size=12).map("".join))
def test_grammar_random_strings(s):
    assert o.int_grammar(s) == _py_int(s)


def test_digit_limit():
    lim = sys.get_int_max_str_digits()
    assert lim == 4300
    for s in ("1" * lim, "1" * (lim + 1), "0" * (lim + 1), "1_" * (lim // 2) + "1", " " * 9000 + "5"):
        assert o.int_grammar(s) == _py_int(s)
    assert o.int_grammar("1" * 50, max_digits=49) is None and o.int_grammar("1" * 50, max_digits=0) == int("1" * 50)


@pytest.mark.needs_reference
def test_restatement_equals_live_reference():
    """Every function against the reference module itself on the adversarial fleet."""
    from oracle.reference_loader import load_reference
    from kubernetes import client as kc
    ref = load_reference()
    assert tuple(ref.GPU_RESOURCE_KEYS) == o.GPU_KEYS
    templates, tid = synth.fleet("g2", 3, 0, 4000)
    nodes = [synth.build_node(templates[t], {"name": f"n{i}"}) for i, t in enumerate(tid)]
    for n in nodes[:1500]:
        assert ref.is_ready(n) == o.ready_of(n)
        assert ref.gpu_capacity(n) == o.gpu_breakdown_of(n)
        assert ref.extract_node_info(n) == o.node_record(n)
    kc.CoreV1Api.set_fleet(nodes)
    try:
        rg, rr = ref.list_gpu_nodes(kc.CoreV1Api())
        og, orr = o.list_gpu_nodes(kc.CoreV1Api())
    finally:
        kc.CoreV1Api.set_fleet(None)
    assert rg == og and rr == orr
    assert all(a is b for a, b in zip(orr, [x for x in og if x["ready"]]))     # same dict objects (:223-225)
