"""The C packer (csrc/fastpack.c) produces exactly the bytes of the Python packer."""
import collections
import time

import numpy as np
import pytest

from k8s_gpu_node_checker_b200 import native, pnl, synth


@pytest.fixture(scope="module", autouse=True)
def built():
    if pnl._fastpack() is None:
        native.build_fastpack()
    assert pnl._fastpack() is not None


def _same(nodes):
    a = pnl.pack_nodes(nodes, use_c=False)
    b = pnl.pack_nodes(nodes, use_c=True)
    assert a.n_nodes == b.n_nodes
    np.testing.assert_array_equal(a.row_ptr, b.row_ptr)
    assert a.rows.tobytes() == b.rows.tobytes()


def test_every_template():
    tmpl = synth.g2_templates()
    _same([synth.build_node(t, {"name": f"n{i}"}) for i, t in enumerate(tmpl)])
    _same([])


def test_fleets_with_payload():
    for kind in ("g1", "g2"):
        tmpl, tid = synth.fleet(kind, 9, 0, 4000)
        _same([synth.build_node(tmpl[t], synth.g1_payload(i, tmpl[t]) if t < 1224 else None) for i, t in enumerate(tid)])


def test_odd_objects():
    from kubernetes import client as kc

    class Weird:
        def __bool__(self):
            raise RuntimeError("no truth")

    class Loud:
        def __str__(self):
            raise RuntimeError("no str")

    caps = collections.OrderedDict([("nvidia.com/gpu", 3), (7, "x"), ("amd.com/gpu", Weird()), ("intel.com/gpu", Loud()),
                                    ("k", b"4"), ("f", 2.5), ("z", 0), ("s", "\ud800"), ("n", None)])
    conds = (kc.V1NodeCondition(type="Ready", status="True"), object(), kc.V1NodeCondition(type=None, status=3),
             kc.V1NodeCondition(type="\ud800", status="x" * 300))
    n = kc.V1Node(status=kc.V1NodeStatus(capacity=caps, conditions=conds))
    _same([n, kc.V1Node(), kc.V1Node(status=kc.V1NodeStatus(capacity={}, conditions=[]))])


def test_it_is_faster():
    tmpl, tid = synth.fleet("g1", 1, 0, 20000)
    objs = [synth.build_node(t) for t in tmpl]
    nodes = [objs[t] for t in tid]
    t0 = time.perf_counter()
    pnl.pack_nodes(nodes, use_c=False)
    t1 = time.perf_counter()
    pnl.pack_nodes(nodes, use_c=True)
    t2 = time.perf_counter()
    print(f"python {20000 / (t1 - t0):.0f} nodes/s, C {20000 / (t2 - t1):.0f} nodes/s")
    assert (t2 - t1) < (t1 - t0)
