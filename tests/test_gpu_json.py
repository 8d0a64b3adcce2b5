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
        if k != 7:      # doc 7 has true/containers as quantities: the device writes "" where the client writes "True"/"{...}";
            #             neither survives int(), the classification above is identical, only the unused text differs
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
