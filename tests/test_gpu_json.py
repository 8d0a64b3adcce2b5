"""Config 5: raw /api/v1/nodes JSON ingested by the GPU structural scan, feeding the same kernels.

Truth is `json.loads` + the client-style adapter (oracle/json_adapter.py) + the reference's functions;
parity at that boundary is unpinned in the reference, so the adapter's rules are the
contract and are written down there.  Comparisons are bit-exact."""
import json

import numpy as np
import pytest
import torch

from k8s_gpu_node_checker_b200 import engine, native, pnl, synth
from oracle import cgn_oracle as o
from oracle import json_adapter as ja
from oracle import pnl_oracle as po
from tests.helpers import assert_same_outputs, expected_outputs, template_truth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def clf():
    c = engine.Classifier(0)
    yield c
    c.close()


def _device_classify(clf, doc: bytes, nodes_cap: int):
    t = torch.frombuffer(bytearray(doc), dtype=torch.uint8).cuda()
    batch, info = clf.json_to_pnl(t, nodes_cap)
    out = engine.DeviceResult.alloc(batch.n_nodes, clf.device, defer_cap=1 << 15)
    clf.classify(batch, out)
    torch.cuda.synchronize()
    counts = out.counts.cpu().numpy()
    return batch, {"flags": out.flags.cpu().numpy()[:batch.n_nodes], "gpu_idx": out.gpu_idx.cpu().numpy().view(np.uint32),
                   "gpu_total": out.gpu_total.cpu().numpy(), "gpu_vals": out.gpu_vals.cpu().numpy(), "counts": counts}


def _oracle_classify(doc: bytes):
    nodes = ja.nodes_from_json(doc)
    p = pnl.pack_nodes(nodes)
    return nodes, p, po.classify(p.rows, p.row_ptr, p.n_nodes)


@pytest.mark.parametrize("style", [0, 1, 2, 4, 7])
def test_json_scan_equals_json_loads(clf, style):
    tmpl, tid = synth.fleet("g2", 21, 0, 2500)
    doc = synth.nodelist_json(tmpl, tid, style_fn=lambda i: style)
    batch, got = _device_classify(clf, doc, 4000)
    nodes, p, ora = _oracle_classify(doc)
    assert batch.n_nodes == len(nodes) == 2500
    assert_same_outputs(got, ora, f"style {style}")
    # and the reference's own view of the same document
    g, r = o.classify_fleet(nodes)
    d = int(got["counts"][2])
    assert int(got["counts"][0]) + sum(1 for n in g if abs(n["gpus"]) >= 10 ** 18 or any(abs(v) >= 10 ** 18 for v in n["gpu_breakdown"].values())) >= len(g) - d


@pytest.mark.parametrize("style", [0, 1, 2, 3])
def test_rows_are_the_packers_rows(clf, style):
    """Without duplicate keys the device-built PNL rows are byte-identical to the host packer's."""
    tmpl, tid = synth.fleet("g1", 23, 0, 2000)
    doc = synth.nodelist_json(tmpl, tid, style_fn=lambda i: style)
    batch, got = _device_classify(clf, doc, 2000)
    nodes, p, ora = _oracle_classify(doc)
    assert_same_outputs(got, ora, f"g1 style {style}")
    assert batch.rows_bytes == p.n_bytes
    assert batch.rows.cpu().numpy()[:p.n_bytes].tobytes() == p.rows[:p.n_bytes].tobytes()
    np.testing.assert_array_equal(batch.row_ptr.cpu().numpy().view(np.uint32)[:2001], p.row_ptr[:2001])


def test_mixed_styles_and_decorations(clf):
    tmpl, tid = synth.fleet("g2", 22, 0, 3000)
    doc = synth.nodelist_json(tmpl, tid, style_fn=lambda i: i % 8)
    batch, got = _device_classify(clf, doc, 3000)
    _, _, ora = _oracle_classify(doc)
    assert_same_outputs(got, ora, "mixed styles")


DOCS = [
    b'{"items":[]}',
    b'{"kind":"NodeList","items":[ ]  }',
    b'{"items":[{}]}',
    b'{"items":[{"status":{"capacity":{"nvidia.com/gpu":"4"},"conditions":[{"type":"Ready","status":"True"}]}}]}',
    b'{"items":[1,null,"x",[{"status":{}}],{"status":null},{"status":{"capacity":{"amd.com/gpu":2}}}]}',
    b'{"other":[{"status":{"capacity":{"nvidia.com/gpu":"9"}}},{"a":[1,2]}],"items":[{"status":{"capacity":{"nvidia.com/gpu":"1"}}}],"after":[{"x":1},{"y":2}]}',
    b'{"items":[{"status":{"capacity":{"nvidia.com/gpu":"7"}}}],"items":[{"status":{"capacity":{"intel.com/gpu":"3"}}},{}]}',
    b'{"items":[{"status":{"capacity":{"nvidia.com/gpu":0,"amd.com/gpu":-0.0,"intel.com/gpu":4.0,"gpu.intel.com/i915":12}}},'
    b'{"status":{"capacity":{"nvidia.com/gpu":true,"amd.com/gpu":false,"intel.com/gpu":null,"gpu.intel.com/i915":{"a":1}}}}]}',
    b'{"items":[{"status":{"capacity":{"nvidia.com\\/gpu":"\\u0034","k\\"ey":"v\\\\","amd.com/gpu":"\\ud835\\udfdc","x":"\\ud800"},'
    b'"conditions":[{"type":"R\\u0065ady","status":"Tru\\u0065"},{"status":"True"},{"type":"Ready"},[],7]}}]}',
    b' \n{ "items" :\t[ { "status" : { "conditions" : [ ] , "capacity" : { } } } , { "st\\u0061tus" : { "c\\u0061pacity" : '
    b'{ "nvidia.com/gpu" : " 8 " } } } ] } \n',
    # non-string scalars go through the client's str(): a JSON true in `status` is "True" (Ready), the integer
    # zero is "0" (kept in gpu_breakdown with value 0) on a node that is a GPU node through another key
    b'{"items":[{"status":{"capacity":{"nvidia.com/gpu":"2"},"conditions":[{"type":"Ready","status":true}]}},'
    b'{"status":{"capacity":{"nvidia.com/gpu":-0,"amd.com/gpu":1,"intel.com/gpu":0},"conditions":[{"type":"Ready","status":false},'
    b'{"type":true,"status":"True"}]}},{"status":{"capacity":{"nvidia.com/gpu":0},"conditions":[{"status":true,"type":"Ready"}]}}]}',
]


@pytest.mark.parametrize("k", range(len(DOCS)))
def test_small_documents(clf, k):
    doc = DOCS[k]
    json.loads(doc)                                        # the document is valid JSON
    batch, got = _device_classify(clf, doc, 16)
    nodes, p, ora = _oracle_classify(doc)
    assert batch.n_nodes == len(nodes)
    if len(nodes):
        assert_same_outputs(got, ora, f"doc {k}")
        if k != 7:      # doc 7 has floats/true/containers as quantities: the device writes "" where the client writes "-0.0"/"True"/"{...}";
            #             none survives int(), the classification above is identical, only the unused text differs
            assert batch.rows.cpu().numpy()[:p.n_bytes].tobytes() == p.rows[:p.n_bytes].tobytes()
    else:
        assert not got["counts"].any()


def test_capacity_errors_are_reported(clf):
    t = torch.frombuffer(bytearray(b'{"nodes":[{}]}'), dtype=torch.uint8).cuda()
    with pytest.raises(native.GncError, match="items"):
        clf.json_to_pnl(t, 8)
    doc = synth.nodelist_json(synth.g1_templates(), np.arange(50))
    t = torch.frombuffer(bytearray(doc), dtype=torch.uint8).cuda()
    with pytest.raises(native.GncError, match="nodes_cap"):
        clf.json_to_pnl(t, 10)
    with pytest.raises(native.GncError, match="rows_cap"):
        clf.json_to_pnl(t, 64, rows_cap=1024)


def test_300k_node_document(clf, golden):
    """A G1 fleet as one ~200 MB NodeList document: same flags, indices, values and counts as the
    reference's per-template truth."""
    tmpl = synth.g2_templates()
    truth = template_truth(golden("templates.json")["records"])
    n = 300_000
    _, tid = synth.fleet("g1", 4, 0, n)
    cache = {}
    parts = []
    for i, t in enumerate(tid):
        t = int(t)
        if t not in cache:
            cache[t] = synth.node_json(tmpl[t], {"name": "node-x"})
        parts.append(cache[t])
    doc = ('{"kind":"NodeList","apiVersion":"v1","metadata":{"resourceVersion":"1"},"items":[' + ",".join(parts) + "]}").encode()
    batch, got = _device_classify(clf, doc, n)
    assert batch.n_nodes == n
    assert_same_outputs(got, expected_outputs(truth, tid), "300k json")


# ---- the paths of the structural-index walker ---------------------------------------------------------------
def _node(cap=None, conds=None, between="", before_status="", name="n"):
    """One element of items, compact, with room for junk between capacity and conditions."""
    st = []
    if cap is not None:
        st.append('"capacity":{' + ",".join(json.dumps(k) + ":" + json.dumps(v) for k, v in cap) + "}")
    if between:
        st.append(between)
    if conds is not None:
        st.append('"conditions":[' + ",".join('{"lastHeartbeatTime":"2025-01-01T00:00:00Z","status":%s,"type":%s,"reason":"r"}'
                                              % (json.dumps(s), json.dumps(t)) for t, s in conds) + "]")
    return '{"metadata":{"name":%s}%s,"status":{%s}}' % (json.dumps(name), before_status, ",".join(st))


def _doc(nodes):
    return ('{"kind":"NodeList","items":[' + ",".join(nodes) + "]}").encode()


def _check_rows(clf, doc, n, want_general=None):
    json.loads(doc)
    batch, got = _device_classify(clf, doc, n)
    nodes, p, ora = _oracle_classify(doc)
    assert batch.n_nodes == len(nodes) == n
    assert_same_outputs(got, ora, "walker paths")
    assert batch.rows_bytes == p.n_bytes
    assert batch.rows.cpu().numpy()[:p.n_bytes].tobytes() == p.rows[:p.n_bytes].tobytes()
    np.testing.assert_array_equal(batch.row_ptr.cpu().numpy().view(np.uint32)[:n + 1], p.row_ptr[:n + 1])
    t = torch.frombuffer(bytearray(doc), dtype=torch.uint8).cuda()
    _, info = clf.json_to_pnl(t, n)
    if want_general is not None:
        assert int(info[5]) == want_general, f"general-parser nodes {int(info[5])}, expected {want_general}"
    return int(info[5])


READY = [("MemoryPressure", "False"), ("Ready", "True")]


def test_plain_documents_take_the_index_walker(clf):
    """Ordinary API output never touches the byte-serial parser; escapes, duplicates and non-string
    quantities in the parts that matter send exactly those nodes to it."""
    tmpl, tid = synth.fleet("g1", 5, 0, 3000)
    doc = synth.nodelist_json(tmpl, tid, style_fn=lambda i: i % 2)          # plain and whitespace-decorated
    t = torch.frombuffer(bytearray(doc), dtype=torch.uint8).cuda()
    _, info = clf.json_to_pnl(t, 3000)
    assert int(info[5]) == 0
    nodes = [_node([("nvidia.com/gpu", "4")], READY) for _ in range(40)]
    nodes[3] = _node([("nvidia.com/gpu", 4)], READY)                         # a number where a string is read
    nodes[7] = nodes[7].replace('"capacity"', '"c\\u0061pacity"')            # an escaped spelling of a key that matters
    nodes[11] = nodes[11].replace('"status":{', '"status":{"capacity":{"amd.com/gpu":"1"},')   # duplicate key: last wins
    nodes[13] = _node([("nvidia.com/gpu", "\\u0038")], READY).replace("\\\\u", "\\u")         # an escape in a quantity
    nodes[17] = _node([("nvidia.com/gpu", "2")], [("Ready", None)])          # null where a string is read
    nodes[19] = nodes[19].replace('"type":"Ready"', '"type":"Ready","type":"Gone"')            # duplicate inside a condition
    _check_rows(clf, _doc(nodes), 40, want_general=6)


def test_rows_too_big_for_notes(clf):
    """More than 16 entries (no notes: the writing pass walks the index again), entries of 255 bytes and more
    (extended rows), and both mixed with ordinary nodes in one warp."""
    many = [(f"example.com/res-{i}", str(i)) for i in range(40)] + [("nvidia.com/gpu", "8")]
    conds20 = [(f"Cond{i}", "False") for i in range(20)] + [("Ready", "True")]
    longv = [("nvidia.com/gpu", "0" * 300 + "7"), ("k" * 400, "1")]
    wide = [(f"x{i}", "1") for i in range(300)] + [("amd.com/gpu", "2")]
    nodes = []
    for i in range(200):
        kind = i % 5
        nodes.append(_node(many, READY) if kind == 0 else _node([("amd.com/gpu", "1")], conds20) if kind == 1
                     else _node(longv, READY) if kind == 2 else _node(wide, READY) if kind == 3
                     else _node([("nvidia.com/gpu", "4")], READY))
    _check_rows(clf, _doc(nodes), 200, want_general=0)


def test_texts_far_apart_and_big_blocks(clf):
    """Capacity and conditions more than a gather span apart (fetched by the owning thread), conditions
    before capacity (notes refused: capacity is written first), and 128 nodes whose rows exceed the CTA's
    staging buffer (written straight to global memory)."""
    junk = '"images":[' + ",".join('{"names":["registry.example/%d"],"sizeBytes":%d}' % (i, i) for i in range(60)) + "]"
    far = [_node([("nvidia.com/gpu", "2")], READY, between=junk) for _ in range(70)]
    _check_rows(clf, _doc(far), 70, want_general=0)
    rev = []
    for i in range(70):
        s = _node([("intel.com/gpu", "1"), ("cpu", "8")], READY)
        a, b = s.index('"capacity"'), s.index(',"conditions"')
        cap, rest = s[a:b], s[b + 1:-2]
        rev.append(s[:a] + rest + "," + cap + "}}")
    _check_rows(clf, _doc(rev), 70, want_general=0)
    fat = [(f"vendor.example/resource-with-a-long-name-{i:02d}", "1" * 40) for i in range(15)] + [("nvidia.com/gpu", "1")]
    _check_rows(clf, _doc([_node(fat, READY) for _ in range(300)]), 300, want_general=0)
    fat2 = [(f"vendor.example/r{i:02d}", "7" * 200) for i in range(12)] + [("nvidia.com/gpu", "1")]      # notes kept, rows of ~3 KB
    _check_rows(clf, _doc([_node(fat2, READY) for _ in range(200)]), 200, want_general=0)


def test_elements_that_are_not_objects_and_odd_values(clf):
    nodes = ['1', '"s"', '[{"status":{"capacity":{"nvidia.com/gpu":"1"}}}]', 'null', '{}', '{"status":7}', '{"status":"x"}',
             '{"status":[]}', '{"status":{"capacity":[],"conditions":{}}}', '{"status":{"capacity":{},"conditions":[1,"a",[2],{}]}}',
             _node([("nvidia.com/gpu", "3")], [("Ready", "True")]),
             '{"status":{"conditions":[{"type":"Ready","status":"True","extra":{"a":[1,2,{"b":"}"}]}}],"capacity":{"amd.com/gpu":"2"}}}',
             '{"spec":{"taints":[{"key":"k","value":null}]},"status":{"capacity":{"":"5","nvidia.com/gpu":""},"conditions":[{"type":"","status":""}]}}']
    _check_rows(clf, _doc(nodes), len(nodes))


def test_backslash_runs_and_long_strings_across_block_boundaries(clf):
    """The lexer's carries: runs of backslashes of every length and parity, quotes after them, and strings of
    hundreds of kilobytes, slid across the 32-byte group, 256-byte chunk and 32 KB tile boundaries by padding."""
    import random
    r = random.Random(5)
    nodes = []
    for i in range(400):
        k = r.choice([1, 2, 3, 15, 16, 17, 31, 32, 33, 63, 64, 65, 127, 128, 255, 256, 257, 300])
        junk = ("\\" * k) + r.choice(['"', '', 'x', '"}]', '\\"']) + ("\\" * r.randint(0, 3))
        pad = " " * r.randint(0, 70)
        ann = json.dumps({"a": junk, "b": ["\\", "\"", junk[::-1]], pad + "k": junk + '"'})
        gpus = str(1 + i % 8)
        nodes.append('{"metadata":{"annotations":%s,"name":%s},%s"status":{"capacity":{"nvidia.com/gpu":"%s","cpu":"8"},'
                     '"conditions":[{"message":%s,"type":"Ready","status":"%s"}]}}'
                     % (ann, json.dumps("n" + junk), pad, gpus, json.dumps(junk), "True" if i % 3 else "False"))
    big = json.dumps(("\\\"" * 40000) + "}{][,:" * 20000)                      # ~360 KB of escapes and brackets in one string
    nodes.insert(200, '{"metadata":{"annotations":{"big":%s}},"status":{"capacity":{"amd.com/gpu":"2"},"conditions":[]}}' % big)
    nodes.insert(300, '{"status":{"capacity":{"intel.com/gpu":%s},"conditions":[{"type":"Ready","status":"True"}]}}' % json.dumps("7" * 70000))
    doc = _doc(nodes)
    for shift in (0, 1, 13, 31):
        d = (b" " * shift) + doc
        json.loads(d)
        batch, got = _device_classify(clf, d, len(nodes) + 8)
        nodes_o, p, ora = _oracle_classify(d)
        assert batch.n_nodes == len(nodes_o) == len(nodes)
        assert_same_outputs(got, ora, f"shift {shift}")
