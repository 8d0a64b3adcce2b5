"""Parity of the CUDA path (through the C ABI) with the reference's outputs -- needs a B200.

Truth comes from tests/golden/templates.json (the reference's own extract_node_info / list_gpu_nodes
per template) expanded by template id, from the C oracle on the same bytes, and -- for whole
documents -- from the reference's `--json` text (by SHA-256) and exit code in tests/golden/fleets.json.
All comparisons are bit-exact (integer / byte work; no tolerance).
"""
import hashlib
import os

import numpy as np
import pytest
import torch

from k8s_gpu_node_checker_b200 import api, engine, native, pnl, synth
from oracle import pnl_oracle as po
from tests.helpers import (F_DEFERRED, F_GPU, F_MALFORMED, KEYS, assert_same_outputs, expected_outputs, key_total,
                           template_truth)

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def clf():
    c = engine.Classifier(0)
    yield c
    c.close()


@pytest.fixture(scope="module")
def g2(golden):
    tmpl = synth.g2_templates()
    rows = synth.template_rows(tmpl)
    recs = golden("templates.json")["records"]
    return tmpl, rows, template_truth(recs), recs


def run_device(clf, p: pnl.PackedNodeList, node_base=0, defer_cap=4096, out_cap=None):
    b = engine.DeviceBatch.from_packed(p, clf.device, node_base)
    out = engine.DeviceResult.alloc(p.n_nodes, clf.device, out_cap=out_cap, defer_cap=defer_cap)
    clf.classify(b, out)
    torch.cuda.synchronize()
    counts = out.counts.cpu().numpy()
    d = min(int(counts[2]), defer_cap)
    defer = out.defer.cpu().numpy().view(engine._DEFER_DT)[:d]
    return {"flags": out.flags.cpu().numpy()[:p.n_nodes], "gpu_idx": out.gpu_idx.cpu().numpy().view(np.uint32),
            "gpu_total": out.gpu_total.cpu().numpy(), "gpu_vals": out.gpu_vals.cpu().numpy(), "counts": counts,
            "defer": defer}


def test_library_is_the_cuda_one(clf):
    assert native.lib().gnc_abi_version() == native.ABI_VERSION
    assert clf.launches == 0


def test_every_template(clf, g2):
    tmpl, rows, truth, recs = g2
    p = pnl.from_rows(rows)
    got = run_device(clf, p)
    assert_same_outputs(got, expected_outputs(truth, np.arange(len(rows))), "templates")
    assert not (got["flags"] & F_MALFORMED).any()
    # deferred (bignum) records: same nodes as the truth, value texts int() to the reference's values
    order = np.argsort(got["defer"]["node"])
    assert list(got["defer"]["node"][order]) == list(np.nonzero(truth["flags"] & F_DEFERRED)[0])
    for rec in got["defer"]:
        bd = {k: int(p.rows[rec["off"][j]:rec["off"][j] + rec["len"][j]].tobytes().decode("utf-8", "surrogatepass"))
              for j, k in enumerate(KEYS) if rec["off"][j] != 0xFFFFFFFF}
        assert bd == recs[int(rec["node"])]["bd"]
    assert clf.launches >= 1


@pytest.mark.parametrize("kind", ["g1", "g2"])
@pytest.mark.parametrize("n", [1, 31, 32, 255, 256, 257, 1000, 4097, 100000])
def test_fleet_sizes(clf, g2, kind, n):
    """Ragged sizes around the tile size; device == reference truth == C oracle."""
    tmpl, rows, truth, _ = g2
    _, tid = synth.fleet(kind, 1, 17, n)
    p = pnl.from_rows([rows[t] for t in tid])
    got = run_device(clf, p, node_base=1000, defer_cap=1 << 15)
    exp = expected_outputs(truth, tid, node_base=1000)
    assert_same_outputs(got, exp, f"{kind} n={n}")
    ora = po.classify(p.rows, p.row_ptr, n, node_base=1000)
    assert_same_outputs(got, ora, f"{kind} n={n} vs C oracle")
    assert sorted(got["defer"]["node"].tolist()) == [d[0] for d in ora["defer"]]


def test_empty_fleet(clf):
    p = pnl.from_rows([])
    got = run_device(clf, p)
    assert not got["counts"].any()
    res = api.classify_packed(p, clf)
    assert (res.total_nodes, res.ready_nodes, res.exit_code) == (0, 0, 2)


def test_repeat_launches_reuse_scratch(clf, g2):
    """The library's scratch (tickets, look-back descriptors, accumulators) re-arms itself."""
    tmpl, rows, truth, _ = g2
    for rep, (seed, n) in enumerate([(3, 5000), (4, 70000), (5, 300), (3, 5000)]):
        _, tid = synth.fleet("g2", seed, 0, n)
        got = run_device(clf, pnl.from_rows([rows[t] for t in tid]))
        assert_same_outputs(got, expected_outputs(truth, tid), f"rep {rep}")


def test_10m_g1_fleet(clf, g2):
    """Config 3 at full size on one GPU: every flag, index, value and count, bit-exact."""
    tmpl, rows, truth, _ = g2
    n = 10_000_000
    _, tid = synth.fleet("g1", 0, 0, n)
    r, rp, total = synth.expand(rows, torch.from_numpy(tid).cuda())
    b = engine.DeviceBatch(r, rp, n, total)
    out = clf.classify(b)
    torch.cuda.synchronize()
    got = {"flags": out.flags.cpu().numpy(), "gpu_idx": out.gpu_idx.cpu().numpy().view(np.uint32),
           "gpu_total": out.gpu_total.cpu().numpy(), "gpu_vals": out.gpu_vals.cpu().numpy(),
           "counts": out.counts.cpu().numpy()}
    exp = expected_outputs(truth, tid)
    assert_same_outputs(got, exp, "10M g1")
    g = int(got["counts"][0])
    assert 0.69 < g / n < 0.71                                   # the 70 % GPU mix
    assert np.all(np.diff(got["gpu_idx"][:g].astype(np.int64)) > 0)   # fleet order kept
    assert int(got["gpu_total"][:g].sum()) == sum(key_total(got["counts"], k) for k in range(4))


def test_1m_adversarial_fleet(clf, g2):
    """Config 4: 1M nodes, half of them adversarial, including bignum and oversize rows."""
    tmpl, rows, truth, recs = g2
    n = 1_000_000
    _, tid = synth.fleet("g2", 0, 0, n)
    r, rp, total = synth.expand(rows, torch.from_numpy(tid).cuda())
    out = engine.DeviceResult.alloc(n, clf.device, defer_cap=1 << 17)
    clf.classify(engine.DeviceBatch(r, rp, n, total), out)
    torch.cuda.synchronize()
    counts = out.counts.cpu().numpy()
    got = {"flags": out.flags.cpu().numpy(), "gpu_idx": out.gpu_idx.cpu().numpy().view(np.uint32),
           "gpu_total": out.gpu_total.cpu().numpy(), "gpu_vals": out.gpu_vals.cpu().numpy(), "counts": counts}
    exp = expected_outputs(truth, tid)
    assert_same_outputs(got, exp, "1M g2")
    d = int(counts[2])
    defer = out.defer.cpu().numpy().view(engine._DEFER_DT)[:d]
    assert sorted(defer["node"].tolist()) == exp["deferred_nodes"].tolist()
    # fold the bignum nodes in on the host and compare with the reference's per-template membership
    fr = engine.finish(got["flags"], got["gpu_idx"], got["gpu_total"], got["gpu_vals"], counts, defer, r.cpu().numpy())
    in_gpu = np.array([x["in_gpu"] for x in recs], dtype=np.int64)
    in_ready = np.array([x["in_ready"] for x in recs], dtype=np.int64)
    np.testing.assert_array_equal(fr.gpu_idx, np.nonzero(in_gpu[tid])[0])
    assert fr.ready_nodes == int(in_ready[tid].sum())
    assert fr.exit_code == 0


@pytest.mark.parametrize("variant", ["t256g2", "t192g2", "t128g2", "t64g2", "t32", "t224", "t128s3", "t64"])
def test_every_kernel_variant(g2, variant, monkeypatch):
    """Each tile shape, on a fleet with rows too large to stage (read straight from global memory)."""
    tmpl, rows, truth, _ = g2
    monkeypatch.setenv("GNC_VARIANT", variant)
    c = engine.Classifier(0)
    try:
        _, tid = synth.fleet("g2", 9, 0, 20000)
        got = run_device(c, pnl.from_rows([rows[t] for t in tid]))
        assert_same_outputs(got, expected_outputs(truth, tid), variant)
    finally:
        c.close()


def test_host_buffer_entry_point(clf, g2):
    """gnc_classify_host (chunked, pipelined copies) == device-resident path, across chunk boundaries."""
    tmpl, rows, truth, _ = g2
    n = (1 << 21) + 12345
    _, tid = synth.fleet("g1", 2, 0, n)
    r, rp, total = synth.expand(rows, torch.from_numpy(tid).cuda())
    p = pnl.PackedNodeList(r.cpu().numpy(), rp.cpu().numpy().view(np.uint32), n)
    out = clf.classify_host(p)
    assert_same_outputs(out, expected_outputs(truth, tid), "host path")
    _, tid = synth.fleet("g2", 2, 0, 50000)
    p = pnl.from_rows([rows[t] for t in tid])
    out = clf.classify_host(p, defer_cap=1 << 14)
    exp = expected_outputs(truth, tid)
    assert_same_outputs(out, exp, "host path g2")
    d = int(out["counts"][2])
    assert sorted(out["defer"]["node"][:d].tolist()) == exp["deferred_nodes"].tolist()


def test_host_path_contains_corrupt_row_ptr_in_later_chunks(g2, monkeypatch):
    """gnc_classify_host hands the kernel a staging buffer minus the chunk's first offset: a row_ptr entry
    that points below the chunk (or past the data) must give malformed rows, not a read outside the buffer.
    Small chunks put the damage in a second and third chunk, at a tile's first row (the tile is then read
    directly from global memory) and in the middle of a staged tile."""
    tmpl, rows, truth, _ = g2
    monkeypatch.setenv("GNC_HOST_CHUNK", "1000")
    n = 5000
    _, tid = synth.fleet("g1", 5, 0, n)
    p = pnl.from_rows([rows[t] for t in tid])
    exp = expected_outputs(truth, tid)
    c = engine.Classifier(0)
    try:
        clean = c.classify_host(p)
        assert_same_outputs(clean, exp, "host path, small chunks")
        rp = p.row_ptr.copy()
        hit = [1000 + 224, 2500, 2600, 3000 + 448 + 17]          # tile starts of chunk 1 and 3, mid-tile in chunk 2
        rp[hit[0]] = 16
        rp[hit[1]] = 0
        rp[hit[2]] = 0xFFFFFF00
        rp[hit[3]] = rp[3000] - 4 if rp[3000] >= 4 else 0         # just below the chunk's first byte
        out = c.classify_host(pnl.PackedNodeList(p.rows, rp, n))
        touched = np.zeros(n, bool)
        for h in hit:
            touched[h - 1:h + 1] = True
        np.testing.assert_array_equal(out["flags"][:n][~touched], exp["flags"][~touched])
        assert all(out["flags"][h - 1] & F_MALFORMED for h in hit)           # the row that now ends before it starts / past the data
        assert int(out["counts"][12]) >= len(hit) and int(out["counts"][3]) == n
        again = c.classify_host(p)                                         # the context is still usable: nothing faulted
        assert_same_outputs(again, exp, "host path after corrupt input")
    finally:
        c.close()


def test_host_path_reports_deferred_drops_once(g2):
    """defer_cap too small on the host path: counts[13] is the number of deferred nodes left out, once."""
    tmpl, rows, truth, _ = g2
    _, tid = synth.fleet("g2", 2, 0, 50000)
    p = pnl.from_rows([rows[t] for t in tid])
    exp = expected_outputs(truth, tid)
    d = int(exp["counts"][2])
    assert d > 40
    c = engine.Classifier(0)
    try:
        with pytest.raises(native.GncError):
            c.classify_host(p, defer_cap=d - 25)
        out = engine.host_buffers(p.n_nodes, d - 25)
        with pytest.raises(native.GncError):
            c.classify_host(p, defer_cap=d - 25, out=out)
        assert int(out["counts"][13]) == 25 and int(out["counts"][2]) == d
    finally:
        c.close()


def test_malformed_rows_match_oracle(clf):
    from tests.test_oracle_pnl import bad_rows
    good, bad = bad_rows()
    rows = []
    for b in bad:
        rows += [good, b] * 40
    p = pnl.from_rows(rows)
    got = run_device(clf, p)
    ora = po.classify(p.rows, p.row_ptr, p.n_nodes)
    assert_same_outputs(got, ora, "malformed")
    assert int(got["counts"][12]) == 40 * len(bad)
    # a row_ptr that runs backwards or past the data is contained, not followed
    rp = p.row_ptr.copy()
    rp[5], rp[9] = rp[7], 0xFFFFFF00
    q = pnl.PackedNodeList(p.rows, rp, p.n_nodes)
    got = run_device(clf, q)
    assert int(got["counts"][12]) >= 40 * len(bad)


def test_extended_and_misaligned_rows(clf):
    """32-bit-length rows and rows that do not start on a 4-byte boundary take the general walk."""
    caps = [(b"cpu", b"4"), (b"amd.com/gpu", b" 2 "), (b"nvidia.com/gpu", b"7")]
    conds = [(b"DiskPressure", b"False"), (b"Ready", b"True")]
    rows = [pnl.encode_raw_row(caps, conds, extended=False), pnl.encode_raw_row(caps, conds, extended=True),
            pnl.encode_raw_row(caps, conds, extended=False)[:-1], pnl.encode_raw_row(caps[:2], conds[:1], extended=False)] * 100
    p = pnl.from_rows(rows)
    got = run_device(clf, p)
    ora = po.classify(p.rows, p.row_ptr, p.n_nodes)
    assert_same_outputs(got, ora, "extended/misaligned")
    assert int(got["counts"][0]) == 400 and int(got["counts"][1]) == 300 and int(got["counts"][12]) == 0


def test_output_capacity_is_respected(clf, g2):
    tmpl, rows, truth, _ = g2
    _, tid = synth.fleet("g1", 0, 0, 5000)
    p = pnl.from_rows([rows[t] for t in tid])
    exp = expected_outputs(truth, tid)
    g = int(exp["counts"][0])
    got = run_device(clf, p, out_cap=g - 7)
    assert int(got["counts"][14]) == 7 and int(got["counts"][0]) == g
    np.testing.assert_array_equal(got["gpu_idx"][:g - 7], exp["gpu_idx"][:g - 7])
    with pytest.raises(native.GncError):
        engine.finish(got["flags"], got["gpu_idx"], got["gpu_total"], got["gpu_vals"], got["counts"], got["defer"], p.rows)


def test_argument_errors(clf):
    p = pnl.from_rows([b"\x00\x00"] * 10)
    b = engine.DeviceBatch.from_packed(p, clf.device)
    b.rows = b.rows[1:]                                             # misaligned
    with pytest.raises(native.GncError, match="aligned"):
        clf.classify(b)


# ---- whole documents through the reference-facing function ------------------------------------
def test_readme_fixture_json(clf, golden):
    from kubernetes import client as kc
    from oracle.make_golden import readme_nodes
    gold = golden("readme_fixture.json")
    kc.CoreV1Api.set_fleet(readme_nodes())
    try:
        g, r = api.list_gpu_nodes(kc.CoreV1Api(), clf)
    finally:
        kc.CoreV1Api.set_fleet(None)
    assert api.summary_json(g, r) + "\n" == gold["json_stdout"]
    assert api.exit_code(g, r) == gold["json_exit"] == 0
    assert all(a is b for a, b in zip(r, [x for x in g if x["ready"]]))


def test_small_fleets_json_and_exit_codes(clf, golden, g2):
    tmpl = g2[0]
    for name, rec in golden("fleets.json")["small"].items():
        nodes = [synth.build_node(tmpl[tid], p) for tid, p in rec["items"]]
        g, r = api.classify_nodes(nodes, clf)
        assert api.summary_json(g, r) + "\n" == rec["json_stdout"], name
        assert api.exit_code(g, r) == rec["exit"], name


@pytest.mark.parametrize("name", ["g1_seed0_n10000", "g2_seed0_n10000", "g1_seed7_n3000"])
def test_10k_documents(clf, golden, name):
    """Configs 2 and 4 at 10k nodes: the JSON document is byte-identical to the reference's."""
    rec = golden("fleets.json")["big"][name]
    templates, tid = synth.fleet(rec["kind"], rec["seed"], 0, rec["n"])
    nodes = [synth.build_node(templates[t], synth.g1_payload(i, templates[t]) if t < 1224 else {"name": f"adv-{i}"})
             for i, t in enumerate(tid)]
    g, r = api.classify_nodes(nodes, clf)
    txt = api.summary_json(g, r) + "\n"
    assert hashlib.sha256(txt.encode()).hexdigest() == rec["sha256"]
    assert api.exit_code(g, r) == rec["exit"]
    assert (len(g), len(r)) == (rec["total_nodes"], rec["ready_nodes"])
