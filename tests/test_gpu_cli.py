"""SURVEY 8f rows 3 and 4 on the GPU: the CLI wrapper runs the reference's own main() (check-gpu-node.py:298-327)
with the CUDA library behind `list_gpu_nodes`, nothing patched away, and the `--json` text comes from the native
writer fed with a DEVICE-produced result (check-gpu-node.py:273-279).  Expected text and exit codes are the
reference's own (tests/golden/readme_fixture.json, fleets.json, written by oracle/make_golden.py).
The reference reaches the GPU box as the byte-compiled copy in oracle/_ref/ (oracle/build_ref.py)."""
import contextlib
import hashlib
import io

import pytest

from k8s_gpu_node_checker_b200 import api, cli, engine, pnl, synth
from oracle.reference_loader import reference_path

pytestmark = pytest.mark.gpu
needs_ref = pytest.mark.skipif(reference_path() is None, reason="no reference script and no oracle/_ref/check_gpu_node.bin on this machine")


def _run(argv, nodes):
    from kubernetes import client as kc
    kc.CoreV1Api.set_fleet(nodes)
    out = io.StringIO()
    try:
        with contextlib.redirect_stdout(out):
            code = cli.main(argv)
    finally:
        kc.CoreV1Api.set_fleet(None)
    return out.getvalue(), code


@pytest.fixture()
def native_writer_calls(monkeypatch):
    calls = []
    real = api.summary_json_from_result

    def spy(nodes, res):
        calls.append(len(nodes))
        return real(nodes, res)
    monkeypatch.setattr(api, "summary_json_from_result", spy)
    return calls


@needs_ref
def test_cli_b200_engine_matches_reference_text_and_exit_codes(golden, native_writer_calls):
    from oracle.make_golden import readme_nodes
    ref = reference_path()
    gold = golden("readme_fixture.json")
    txt, code = _run(["--reference", ref, "--json"], readme_nodes())
    assert (txt, code) == (gold["json_stdout"], gold["json_exit"])
    assert native_writer_calls == [2]                                  # the --json text came from the native writer
    txt, code = _run(["--reference=" + ref, "--engine", "b200"], readme_nodes())
    assert (txt, code) == (gold["table_stdout"], gold["table_exit"])
    assert native_writer_calls == [2]                                  # table mode does not call it
    g2 = synth.g2_templates()
    seen = set()
    for name, rec in golden("fleets.json")["small"].items():
        nodes = [synth.build_node(g2[tid], p) for tid, p in rec["items"]]
        txt, code = _run(["--reference", ref, "--json", "--slack-only-on-error"], nodes)
        assert (txt, code) == (rec["json_stdout"], rec["exit"]), name
        seen.add(code)
    assert {0, 2, 3} <= seen                                           # check-gpu-node.py:289-293, all three


@needs_ref
@pytest.mark.parametrize("name", ["g1_seed0_n10000", "g2_seed0_n10000"])
def test_cli_10k_documents(golden, name, native_writer_calls):
    rec = golden("fleets.json")["big"][name]
    templates, tid = synth.fleet(rec["kind"], rec["seed"], 0, rec["n"])
    nodes = [synth.build_node(templates[t], synth.g1_payload(i, templates[t]) if t < 1224 else {"name": f"adv-{i}"})
             for i, t in enumerate(tid)]
    txt, code = _run(["--reference", reference_path(), "--json"], nodes)
    assert hashlib.sha256(txt.encode("utf-8", "surrogatepass")).hexdigest() == rec["sha256"]
    assert code == rec["exit"] and native_writer_calls == [rec["n"]]


def test_native_json_writer_on_device_results(golden):
    """summary_json_from_result fed by the CUDA path (not by the oracle's arrays): README fixture, the small
    fleets (bignum nodes included) and a 10k-node document, against the reference's text."""
    from oracle.make_golden import readme_nodes
    clf = engine.Classifier(0)
    try:
        nodes = readme_nodes()
        assert api.summary_json_from_result(*api.classify_for_summary(nodes, clf)) + "\n" == golden("readme_fixture.json")["json_stdout"]
        g2 = synth.g2_templates()
        for name, rec in golden("fleets.json")["small"].items():
            nodes = [synth.build_node(g2[tid], p) for tid, p in rec["items"]]
            nodes, res = api.classify_for_summary(nodes, clf)
            assert api.summary_json_from_result(nodes, res) + "\n" == rec["json_stdout"], name
            assert res.exit_code == rec["exit"], name
            g, r = api.build_records(nodes, res)
            assert (g, r) == api.build_records(nodes, res, use_c=False)            # the C record builder == the Python one
        rec = golden("fleets.json")["big"]["g2_seed0_n10000"]
        templates, tid = synth.fleet(rec["kind"], rec["seed"], 0, rec["n"])
        nodes = [synth.build_node(templates[t], synth.g1_payload(i, templates[t]) if t < 1224 else {"name": f"adv-{i}"})
                 for i, t in enumerate(tid)]
        txt = api.summary_json_from_result(*api.classify_for_summary(nodes, clf)) + "\n"
        assert hashlib.sha256(txt.encode("utf-8", "surrogatepass")).hexdigest() == rec["sha256"]
    finally:
        clf.close()
