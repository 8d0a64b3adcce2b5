"""The C restatement over PNL-v2 bytes (oracle/pnl_oracle.c) and the packer, held to the reference's
own outputs (tests/golden/templates.json, fleets.json)."""
import sys

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from oracle import cgn_oracle as o
from oracle import pnl_oracle as po
from k8s_gpu_node_checker_b200 import pnl, synth
from tests.helpers import (F_DEFERRED, F_GPU, F_MALFORMED, KEYS, assert_same_outputs, expected_outputs, key_total,
                           template_truth)


@pytest.fixture(scope="module")
def g2(golden):
    tmpl = synth.g2_templates()
    rows = synth.template_rows(tmpl)
    truth = template_truth(golden("templates.json")["records"])
    return tmpl, rows, truth


def test_every_template_row(g2):
    """One node per template, in template order: flags, values, counts == reference truth."""
    tmpl, rows, truth = g2
    p = pnl.from_rows(rows)
    got = po.classify(p.rows, p.row_ptr, p.n_nodes)
    exp = expected_outputs(truth, np.arange(len(rows)))
    assert_same_outputs(got, exp, "templates")
    assert not (got["flags"] & F_MALFORMED).any()
    assert int(got["counts"][2]) > 0                      # the >18-digit cases really are there


def test_deferred_records_point_at_value_text(g2, golden):
    """Bignum escape: the record's (offset, length) are the value texts; int() of them == reference."""
    tmpl, rows, truth = g2
    recs = golden("templates.json")["records"]
    p = pnl.from_rows(rows)
    got = po.classify(p.rows, p.row_ptr, p.n_nodes)
    assert [d[0] for d in got["defer"]] == list(np.nonzero(truth["flags"] & F_DEFERRED)[0])
    for node, offs, lens in got["defer"]:
        bd = {}
        for k, key in enumerate(KEYS):
            if offs[k] != 0xFFFFFFFF:
                bd[key] = int(p.rows[offs[k]:offs[k] + lens[k]].tobytes().decode("utf-8", "surrogatepass"))
        assert bd == recs[node]["bd"]


@pytest.mark.parametrize("kind,seed,n", [("g1", 0, 10000), ("g2", 0, 10000), ("g1", 7, 3000)])
def test_fleet_counts_match_reference_main(g2, golden, kind, seed, n):
    """Fleet-level: G and R equal total_nodes / ready_nodes the reference's main() printed."""
    tmpl, rows, truth = g2
    rec = golden("fleets.json")["big"][f"{kind}_seed{seed}_n{n}"]
    _, tid = synth.fleet(kind, seed, 0, n)
    p = pnl.from_rows([rows[t] for t in tid])
    got = po.classify(p.rows, p.row_ptr, n, node_base=5)
    assert_same_outputs(got, expected_outputs(truth, tid, node_base=5), rec["kind"])
    # deferred nodes resolved with Python ints complete the reference's totals
    extra_g = extra_r = 0
    recs = golden("templates.json")["records"]
    for node, _, _ in got["defer"]:
        r = recs[tid[node - 5]]
        extra_g += r["in_gpu"]
        extra_r += r["in_ready"]
    assert int(got["counts"][0]) + extra_g == rec["total_nodes"]
    assert int(got["counts"][1]) + extra_r == rec["ready_nodes"]


def test_spec_rows_equal_packed_objects():
    """synth.spec_row (no client objects) == the packer on the built V1Node, for every template."""
    t = synth.g2_templates()
    assert synth.spec_rows(t) == synth.template_rows(t)
    dup = {"capacity": [["nvidia.com/gpu", "1"], ["cpu", "4"], ["nvidia.com/gpu", "8"]], "conditions": [["plain:Ready", "True"]]}
    assert synth.spec_row(dup) == pnl.node_row(synth.build_node(dup))


def test_expand_equals_row_concatenation(g2):
    """synth.expand (torch gather) builds the same bytes as joining packed rows."""
    tmpl, rows, _ = g2
    _, tid = synth.fleet("g2", 11, 100, 3000)
    r, rp, total = synth.expand(rows, tid)
    p = pnl.from_rows([rows[t] for t in tid])
    assert total == p.n_bytes
    assert r.numpy()[:total].tobytes() == p.rows[:total].tobytes()
    np.testing.assert_array_equal(rp.numpy().view(np.uint32)[:len(tid) + 1], p.row_ptr[:len(tid) + 1])


def test_key_totals_exact_beyond_int64(g2):
    """Per-key totals are carried as 128-bit lo/hi pairs: 30 nodes of 9e17 GPUs overflow int64, not these."""
    row = pnl.encode_row([("nvidia.com/gpu", "9" * 18)], [("Ready", "True")])
    p = pnl.from_rows([row] * 30)
    got = po.classify(p.rows, p.row_ptr, 30)
    assert key_total(got["counts"], 0) == 30 * int("9" * 18)
    row = pnl.encode_row([("amd.com/gpu", "-5"), ("nvidia.com/gpu", "7")], None)
    got = po.classify(*(lambda q: (q.rows, q.row_ptr, 4))(pnl.from_rows([row] * 4)))
    assert key_total(got["counts"], 1) == -20 and key_total(got["counts"], 0) == 28


def test_pack_unpack_roundtrip(g2):
    tmpl, rows, _ = g2
    for t, row in zip(tmpl, rows):
        caps, conds = pnl.decode_row(row)
        if t.get("status", True) is False:
            assert (caps, conds) == ([], [])
            continue
        want = [(k.encode("utf-8", "surrogatepass"), pnl.marshal_quantity(v)) for k, v in dict(map(tuple, t["capacity"] or [])).items()]   # node.status.capacity is a dict
        assert caps == want
        wc = [(a if isinstance(a, str) else "", b if isinstance(b, str) else "") for a, b in (t["conditions"] or [])
              if not (isinstance(a, str) and a.startswith("plain:"))]
        assert conds == [(a.encode(), b.encode()) for a, b in wc]


def bad_rows():
    """Rows whose own lengths do not fit them (each 4-byte padded, as the format requires)."""
    good = pnl.encode_row([("nvidia.com/gpu", "4")], [("Ready", "True")])
    pad = lambda b: b + b"\0" * (-len(b) % 4)   # noqa: E731
    return good, [
        b"\x01",                                         # one byte: no header
        pad(b"\x05\x05\x01\x01"),                        # table longer than the row
        pad(b"\x01\x00\x0e\x09nvidia.com/gpu4"),          # value length overruns the row
        pad(b"\x01\x01\x0e\x01\x05\x09nvidia.com/gpu4ReadyTrue"),   # condition overruns; capacity before it counts
        pad(b"\xff\xff\0\0\x01\0\0"),                   # extended header cut short
        pad(b"\xff\xff\0\0" + b"\xff\xff\xff\x7f" * 2),   # extended counts far beyond the row
        pad(b"\xff\xff\0\0\x01\0\0\0\0\0\0\0\xff\xff\xff\xff\x01\0\0\0k"),   # 4 GiB key length
        good[:-8],                                       # text cut short
    ]


def test_malformed_rows_are_contained():
    """Truncated / lying lengths never read outside the row; the row is flagged, earlier entries count."""
    good, bad = bad_rows()
    rows = []
    for b in bad:
        rows += [good, b]
    p = pnl.from_rows(rows + [good])
    got = po.classify(p.rows, p.row_ptr, p.n_nodes)
    f = got["flags"]
    assert all(f[i] & F_GPU for i in range(0, len(f), 2))
    assert all(f[i] & F_MALFORMED for i in range(1, len(f), 2))
    assert int(got["counts"][12]) == len(bad)
    assert f[2 * 3 + 1] & F_GPU and not f[2 * 3 + 1] & 2     # capacity entry kept, overrunning Ready not seen
    assert f[2 * 7 + 1] & F_GPU and not f[2 * 7 + 1] & 2     # likewise when the text is cut short
    assert int(got["counts"][0]) == len(bad) + 1 + 2


def test_extended_rows_equal_short_rows():
    """The 32-bit-length form classifies exactly like the 8-bit form of the same entries."""
    caps = [(b"cpu", b"4"), (b"amd.com/gpu", b" 2 "), (b"nvidia.com/gpu", b"1" * 23)]
    conds = [(b"DiskPressure", b"False"), (b"Ready", b"True")]
    a = pnl.from_rows([pnl.encode_raw_row(caps, conds, extended=False), pnl.encode_raw_row(caps[:2], conds, extended=False)])
    b = pnl.from_rows([pnl.encode_raw_row(caps, conds, extended=True), pnl.encode_raw_row(caps[:2], conds, extended=True)])
    ra, rb = po.classify(a.rows, a.row_ptr, 2), po.classify(b.rows, b.row_ptr, 2)
    np.testing.assert_array_equal(ra["flags"], rb["flags"])
    np.testing.assert_array_equal(ra["gpu_vals"], rb["gpu_vals"])
    assert ra["flags"][0] & F_DEFERRED and ra["flags"][1] & F_GPU
    assert pnl.decode_row(b.row(0)) == (caps, conds)


_chars = st.sampled_from(list("0123456789_+- \t\n\x0b\x0c\r\x1c\x00kKi.e") +
                         ["٤", "４", "\xa0", " ", "\x85", "\U0001d7dc", "²", "\ud800", "​"])


@settings(max_examples=3000, deadline=None)
@given(st.lists(_chars, max_size=24).map("".join))
def test_c_quantity_parser_equals_int(s):
    try:
        want = int(s)
    except ValueError:
        want = None
    status, v = po.parse_quantity(s.encode("utf-8", "surrogatepass"))
    if s == "" or want is None:
        assert status == 0
    elif abs(want) >= 10 ** 18:
        assert status == 2
    else:
        assert (status, v) == (1, want)


def test_c_quantity_parser_bad_utf8_and_limits():
    for raw in (b"\xff4", b"\xc0\xb4", b"4\xe2\x80", b"\xf4\x90\x80\x804", b"\xed\xa0\x80"):
        assert po.parse_quantity(raw)[0] == 0
    lim = sys.get_int_max_str_digits()
    assert po.parse_quantity(b"1" * lim)[0] == 2 and po.parse_quantity(b"1" * (lim + 1))[0] == 0
    assert po.parse_quantity(b"0" * (lim + 1))[0] == 0 and po.parse_quantity(b"0" * lim) == (1, 0)
    assert po.parse_quantity(b"1" * 30, max_digits=0)[0] == 2
    assert o.int_grammar("0" * 22 + "7") == 7 and po.parse_quantity(b"0" * 22 + b"7") == (1, 7)
