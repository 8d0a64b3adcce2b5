"""The CLI wrapper keeps the reference's command line: with the path replaced, its own main() prints the
same text and returns the same exit codes.  Runs here only (needs /root/reference); the GPU classifier is
stood in for by the oracle so that the wrapper's plumbing is what is under test."""
import contextlib
import io

import pytest

from k8s_gpu_node_checker_b200 import api, cli, synth
from oracle import cgn_oracle
from oracle.reference_loader import reference_path

pytestmark = pytest.mark.needs_reference
REFERENCE_SCRIPT = reference_path()       # the script here; the byte-compiled copy in oracle/_ref on the GPU box


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


def _oracle_classifier(monkeypatch):
    """No GPU in this container: the device's decisions are stood in for by the C oracle (same output arrays), so
    that everything above them -- the seam, the record builder, the native --json writer, the wrapper -- runs."""
    import numpy as np
    from k8s_gpu_node_checker_b200 import engine
    from oracle import pnl_oracle as po

    def classify_packed(p, clf=None):
        o = po.classify(p.rows, p.row_ptr, p.n_nodes)
        defer = np.array([(n, offs, lens) for n, offs, lens in o["defer"]], dtype=engine._DEFER_DT)
        return engine.finish(o["flags"], o["gpu_idx"], o["gpu_total"], o["gpu_vals"], o["counts"], defer, p.rows)
    monkeypatch.setattr(api, "classify_packed", classify_packed)


def test_cli_matches_reference(monkeypatch, golden):
    from oracle.make_golden import readme_nodes
    _oracle_classifier(monkeypatch)
    calls = []
    real = api.summary_json_from_result
    monkeypatch.setattr(api, "summary_json_from_result", lambda n, r: (calls.append(1), real(n, r))[1])
    gold = golden("readme_fixture.json")
    txt, code = _run(["--reference", REFERENCE_SCRIPT, "--json"], readme_nodes())
    assert (txt, code) == (gold["json_stdout"], gold["json_exit"])
    txt, code = _run(["--reference=" + REFERENCE_SCRIPT], readme_nodes())
    assert (txt, code) == (gold["table_stdout"], gold["table_exit"])
    assert calls == [1]                                                       # --json went through the native writer, the table did not
    txt, code = _run(["--engine", "reference", "--reference", REFERENCE_SCRIPT, "--json"], readme_nodes())
    assert (txt, code) == (gold["json_stdout"], gold["json_exit"])
    assert calls == [1]
    g2 = synth.g2_templates()
    for name, rec in golden("fleets.json")["small"].items():
        nodes = [synth.build_node(g2[tid], p) for tid, p in rec["items"]]
        txt, code = _run(["--reference", REFERENCE_SCRIPT, "--json", "--slack-only-on-error"], nodes)
        assert (txt, code) == (rec["json_stdout"], rec["exit"]), name


def test_cli_argument_errors(capsys, monkeypatch):
    monkeypatch.delenv("GNC_REFERENCE_SCRIPT", raising=False)
    assert cli.main(["--json"]) == 1              # without --reference / GNC_REFERENCE_SCRIPT the wrapper refuses
    assert "--reference" in capsys.readouterr().err
    assert cli.main(["--reference", REFERENCE_SCRIPT, "--engine", "tpu"]) == 1
