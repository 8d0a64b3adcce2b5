"""Randomised NodeList documents through the GPU JSON scan, against json.loads + the adapter + the reference's
functions (oracle/json_adapter.py).  The generator writes JSON text directly (so that duplicate keys, escaped
spellings, odd whitespace and values of the wrong type can appear anywhere), seeded: a failure reproduces."""
import json
import random

import numpy as np
import pytest
import torch

from k8s_gpu_node_checker_b200 import engine, synth
from tests.helpers import assert_same_outputs
from tests.test_gpu_json import _device_classify, _oracle_classify

pytestmark = pytest.mark.gpu

GPU_KEYS = ["nvidia.com/gpu", "amd.com/gpu", "gpu.intel.com/i915", "intel.com/gpu"]
OTHER_KEYS = ["cpu", "memory", "pods", "nvidia.com/gpus", "amd.com/gp", "intel.com/gpu ", "Nvidia.com/gpu", "", "status", "type",
              "capacity", "conditions", "k\"q", "back\\slash", "br{ace", "br]ack", "unié", "\U0001d7dc"]
STRINGS = [q for q in synth.QUANTITY_CASES if isinstance(q, str) and len(q) < 64 and "\ud800" not in q]


@pytest.fixture(scope="module")
def clf():
    c = engine.Classifier(0)
    yield c
    c.close()


class Gen:
    def __init__(self, seed):
        self.r = random.Random(seed)

    def ws(self):
        return self.r.choice(["", "", "", " ", "\n", "\t ", " \r\n "])

    def string(self, s):
        """A JSON string literal for s: plain, ASCII-escaped, or every character as \\uXXXX."""
        k = self.r.random()
        if k < 0.70:
            return json.dumps(s, ensure_ascii=False)
        if k < 0.85:
            return json.dumps(s, ensure_ascii=True)
        out = []
        for ch in s:
            cp = ord(ch)
            if cp > 0xFFFF:
                cp -= 0x10000
                out.append("\\u%04x\\u%04x" % (0xD800 + (cp >> 10), 0xDC00 + (cp & 0x3FF)))
            else:
                out.append("\\u%04X" % cp if self.r.random() < 0.5 else "\\u%04x" % cp)
        return '"' + "".join(out) + '"'

    def junk(self, depth=0):
        k = self.r.random()
        if depth > 2 or k < 0.35:
            return self.r.choice(['1', '-2.5e3', 'true', 'false', 'null', '"x"', '"}"', '"]["', '"a\\"b"', '"\\\\"', '""', '0'])
        if k < 0.65:
            return "[" + self.ws() + ("," + self.ws()).join(self.junk(depth + 1) for _ in range(self.r.randint(0, 3))) + self.ws() + "]"
        return self.obj([(self.r.choice(OTHER_KEYS), self.junk(depth + 1)) for _ in range(self.r.randint(0, 3))])

    def obj(self, members):
        return "{" + self.ws() + ("," + self.ws()).join(self.string(k) + self.ws() + ":" + self.ws() + v for k, v in members) + self.ws() + "}"

    def quantity(self):
        k = self.r.random()
        if k < 0.55:
            return self.string(self.r.choice(["1", "2", "4", "8", "0", "16"]))
        if k < 0.85:
            return self.string(self.r.choice(STRINGS))
        return self.r.choice(['4', '0', '-0.0', '8.0', '-2', 'true', 'false', 'null', '[1]', '{"a":"4"}', '1e2', '12345678901234567890123'])

    def capacity(self):
        k = self.r.random()
        if k < 0.08:
            return self.r.choice(['null', '[]', '"x"', '7', '[{"nvidia.com/gpu":"4"}]'])
        members = []
        for _ in range(self.r.randint(0, 6)):
            key = self.r.choice(GPU_KEYS) if self.r.random() < 0.5 else self.r.choice(OTHER_KEYS)
            members.append((key, self.quantity()))
        if self.r.random() < 0.01:
            members += [(f"pad{i}", '"1"') for i in range(self.r.choice([20, 260]))]
        return self.obj(members)

    def condition(self):
        k = self.r.random()
        if k < 0.08:
            return self.r.choice(['1', '"Ready"', 'null', '["Ready","True"]', '[]', 'true'])
        members = []
        t = self.r.choice(["Ready", "Ready", "Ready", "ready", "MemoryPressure", "Ready ", "", "Réady"])
        s = self.r.choice(["True", "True", "False", "Unknown", "true", "TRUE", "", "True "])
        if self.r.random() < 0.9:
            members.append(("type", self.string(t) if self.r.random() < 0.93 else self.r.choice(['null', '1', '["Ready"]'])))
        if self.r.random() < 0.9:
            members.append(("status", self.string(s) if self.r.random() < 0.93 else self.r.choice(['null', 'true', '{}'])))
        for _ in range(self.r.randint(0, 3)):
            members.append((self.r.choice(["reason", "message", "lastHeartbeatTime", "extra"]), self.junk(1)))
        if self.r.random() < 0.08:                       # a duplicate: the later one wins
            members.append((self.r.choice(["type", "status"]), self.string(self.r.choice(["Ready", "True", "False"]))))
        self.r.shuffle(members)
        return self.obj(members)

    def conditions(self):
        if self.r.random() < 0.06:
            return self.r.choice(['null', '{}', '"x"', '{"type":"Ready","status":"True"}'])
        return "[" + self.ws() + ("," + self.ws()).join(self.condition() for _ in range(self.r.randint(0, 5))) + self.ws() + "]"

    def status(self):
        if self.r.random() < 0.06:
            return self.r.choice(['null', '[]', '"Ready"', '0', '[{"capacity":{"nvidia.com/gpu":"1"}}]'])
        members = []
        if self.r.random() < 0.9:
            members.append(("capacity", self.capacity()))
        if self.r.random() < 0.9:
            members.append(("conditions", self.conditions()))
        if self.r.random() < 0.5:
            members.append(("allocatable", self.capacity()))
        for _ in range(self.r.randint(0, 3)):
            members.append((self.r.choice(["nodeInfo", "images", "addresses", "capacityX", "Conditions"]), self.junk()))
        if self.r.random() < 0.07:
            members.append(("capacity", self.capacity()))
        if self.r.random() < 0.07:
            members.append(("conditions", self.conditions()))
        self.r.shuffle(members)
        return self.obj(members)

    def node(self):
        if self.r.random() < 0.04:
            return self.r.choice(['1', 'null', '"node"', '[]', '[{"status":{"capacity":{"nvidia.com/gpu":"1"}}}]', '{}', 'true'])
        members = [("metadata", self.junk()), ("spec", self.junk())]
        if self.r.random() < 0.95:
            members.append(("status", self.status()))
        if self.r.random() < 0.05:
            members.append(("status", self.status()))
        if self.r.random() < 0.3:
            members.append((self.r.choice(["kind", "apiVersion", "Status", "statusX"]), self.junk()))
        self.r.shuffle(members)
        return self.obj(members)

    def document(self, n):
        members = [("kind", '"NodeList"'), ("metadata", self.junk()), ("items", "[" + self.ws() + ("," + self.ws()).join(self.node() for _ in range(n)) + self.ws() + "]")]
        if self.r.random() < 0.2:
            members.append(("other", "[" + self.node() + "," + self.node() + "]"))
        if self.r.random() < 0.1:
            members.insert(0, ("items", "[" + self.node() + "]"))          # an earlier "items": the later one wins
        if self.r.random() < 0.5:
            self.r.shuffle(members)
            # keep the winning "items" last among the items members: shuffle may reorder them, so re-derive n from json.loads below
        return (self.ws() + self.obj(members) + self.ws()).encode()


@pytest.mark.parametrize("seed", range(12))
def test_random_documents(clf, seed):
    g = Gen(1000 + seed)
    for k in range(40):
        n = g.r.randint(0, 40)
        doc = g.document(n)
        root = json.loads(doc)                              # valid JSON by construction
        want = len(root["items"])
        batch, got = _device_classify(clf, doc, 64)
        nodes, p, ora = _oracle_classify(doc)
        assert batch.n_nodes == len(nodes) == want, f"seed {seed} doc {k}"
        if want:
            assert_same_outputs(got, ora, f"seed {seed} doc {k}: {doc[:200]!r}")
        else:
            assert not got["counts"].any()


def test_many_random_nodes_in_one_document(clf):
    """Whole warps of unlike nodes: the index walker, its re-walk and the general parser side by side."""
    g = Gen(77)
    doc = g.document(5000)
    n = len(json.loads(doc)["items"])
    batch, got = _device_classify(clf, doc, 8192)
    nodes, p, ora = _oracle_classify(doc)
    assert batch.n_nodes == n
    assert_same_outputs(got, ora, "5000 random nodes")
    t = torch.frombuffer(bytearray(doc), dtype=torch.uint8).cuda()
    _, info = clf.json_to_pnl(t, 8192)
    assert 0 < int(info[5]) < n                             # both parsers were exercised
