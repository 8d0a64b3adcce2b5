"""Seeded synthetic fleets: G1 mixed fleet, G2 adversarial fleet.

A fleet is N draws from a finite list of node *templates*; node i picks its template from
`splitmix64(seed * 0x100000001B3 + i)`, so any slice of any fleet can be regenerated anywhere
(host, GPU box, another rank) without shipping it.  A template is a JSON-able node spec:

    {"capacity": [[key, value], ...] | None,      value: str | int | float | bool | None
     "conditions": [[type, status], ...] | None,  a pair prefixed "plain:" is a non-V1NodeCondition
     "status": False                              -> node.status is None
     "name"/"labels"/"taints"/"metadata": ...     host-side payload (optional)}

`expand` turns (template rows, template ids) into PNL-v2 columns with torch ops, on whatever
device the ids live on, so a 10M-node fleet is built on the GPU in HBM and a 10k-node one on
the CPU by the same code.
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional, Sequence

import numpy as np

from . import pnl

MASK64 = (1 << 64) - 1


def splitmix64(x: np.ndarray) -> np.ndarray:
    """Vectorised splitmix64 finaliser over uint64 (wraps mod 2^64)."""
    with np.errstate(over="ignore"):
        z = (x.astype(np.uint64) + np.uint64(0x9E3779B97F4A7C15))
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def node_hash(seed: int, start: int, n: int) -> np.ndarray:
    base = (seed * 0x100000001B3) & MASK64
    with np.errstate(over="ignore"):
        return splitmix64(np.uint64(base) + np.arange(start, start + n, dtype=np.uint64))


# ----------------------------------------------------------------------------------------------
# G1: configs 2 and 3.
# ----------------------------------------------------------------------------------------------
_CPU = ("32", "64", "96", "128")
_EPH = ("203070420Ki", "939467288Ki")
_MEM = ("263856340Ki", "527951004Ki", "1056283244Ki")
_GPU_OPTS = [None] + [(k, str(c)) for k in pnl.GPU_RESOURCE_KEYS for c in (1, 2, 4, 8)]   # 17
_READY = ("True", "False", "Unknown")
_PRESSURE = [["NetworkUnavailable", "False"], ["MemoryPressure", "False"],
             ["DiskPressure", "False"], ["PIDPressure", "False"]]


def _g1_template(cpu: int, eph: int, mem: int, gpu: int, ready: int) -> Dict[str, Any]:
    cap = {"cpu": _CPU[cpu], "ephemeral-storage": _EPH[eph], "hugepages-1Gi": "0",
           "hugepages-2Mi": "0", "memory": _MEM[mem], "pods": "110"}
    if _GPU_OPTS[gpu]:
        cap[_GPU_OPTS[gpu][0]] = _GPU_OPTS[gpu][1]
    # alphabetical key order, as the API server's JSON map encoding gives
    return {"capacity": [[k, cap[k]] for k in sorted(cap)],
            "conditions": _PRESSURE + [["Ready", _READY[ready]]]}


def g1_templates() -> List[Dict[str, Any]]:
    """All 4*2*3*17*3 = 1224 G1 node templates; index = g1_tid()."""
    return [_g1_template(c, e, m, g, r)
            for c in range(4) for e in range(2) for m in range(3) for g in range(17) for r in range(3)]


def g1_tid(seed: int, start: int, n: int) -> np.ndarray:
    """Template id per node: 30 % no GPU key, else 60/20/10/10 nvidia/amd/i915/intel with
    1/2/4/8 GPUs; Ready 80 % True, 10 % False, 10 % Unknown."""
    h = node_hash(seed, start, n)
    u = lambda sh, bits: (h >> np.uint64(sh)) & np.uint64((1 << bits) - 1)   # noqa: E731
    has_gpu = (u(0, 16) % np.uint64(100)) >= 30
    vend = u(16, 16) % np.uint64(10)
    key = np.select([vend < 6, vend < 8, vend < 9], [0, 1, 2], 3).astype(np.int64)
    cnt = u(32, 2).astype(np.int64)
    gpu = np.where(has_gpu, 1 + key * 4 + cnt, 0)
    rd = u(34, 16) % np.uint64(10)
    ready = np.select([rd < 8, rd < 9], [0, 1], 2).astype(np.int64)
    cpu = u(50, 2).astype(np.int64)
    eph = u(52, 1).astype(np.int64)
    mem = (u(53, 11) % np.uint64(3)).astype(np.int64)
    return ((((cpu * 2 + eph) * 3 + mem) * 17 + gpu) * 3 + ready).astype(np.int32)


def g1_payload(i: int, tmpl: Dict[str, Any]) -> Dict[str, Any]:
    """Host-side payload of G1 node i: name, 4 labels, one taint on nodes with a GPU key."""
    gpu_key = next((k for k, _ in tmpl["capacity"] if k in pnl.GPU_RESOURCE_KEYS), None)
    out = {"name": f"node-{i:08d}",
           "labels": {"kubernetes.io/hostname": f"node-{i:08d}", "kubernetes.io/arch": "amd64",
                      "node.kubernetes.io/instance-type": ("p5.48xlarge" if gpu_key else "m6i.8xlarge"),
                      "topology.kubernetes.io/zone": f"zone-{i % 3}"}}
    if gpu_key:
        out["taints"] = [{"key": gpu_key, "value": "present", "effect": "NoSchedule",
                          "time_added": "2025-01-01T00:00:00Z"}]
    return out


# ----------------------------------------------------------------------------------------------
# G2: config 4 (adversarial).
# ----------------------------------------------------------------------------------------------
QUANTITY_CASES: List[Any] = [
    "4", "0", "-0", "00012", "+4", "-1", " 4 ", "\t4\n", "\x0c4\x0b", "\xa04", " 4 ",
    "\x854", "4 ", "4　", " 4", "1_0", "0_0", "1_2_3", "٤", "４",
    "١٢", "1２", "-٤", "٤_٥", "\U0001d7dc", "1" * 23, "-" + "1" * 23,
    "9" * 18, "-" + "9" * 18, "1" + "0" * 18, "9223372036854775807", "9223372036854775808",
    "-9223372036854775808", "0" * 22 + "7", "1_" * 11 + "1", "1" * 4300, "1" * 4301, "0" * 4301,
    " " * 300 + "8", "1_" * 2200 + "1",
    "1k", "2Ki", "4.0", "1e3", "0x10", "0b1", "0o7", "1L", "1__0", "_1", "1_", "--1", "+ 4", "4 4",
    "1\x00", "\x004", "\x1c4", "4\x1f", "\x7f4", "​4", "﻿4", "᠎4", "−4", "1,000",
    "+", "-", "_", " ", "", "+_1", "-_4", "+-4", "4+", "²", "①", "\ud800", "٤k",
    None, 4, 0, 8.0, True, False, -2, 10 ** 25,
]

KEY_DECOYS = ["nvidia.com/gpu ", "NVIDIA.com/gpu", "nvidia.com/gpv", "nvidia.com/gp", "nvidia.com/gpux",
              "amd.com/gpU", "xmd.com/gpu", "gpu.intel.com/i91", "gpu.intel.com/i916", "intel.com/gpu\x00",
              "hugepages-1Gi", "example.com/foo", "", "n", "nvidia.com/gpú", "k" * 255, "k" * 300]

CONDITION_CASES: List[Any] = [
    None, [], [["Ready", "True"]], [["Ready", "False"]], [["Ready", "Unknown"]],
    [["Ready", "False"], ["Ready", "True"]], [["Ready", "True"], ["Ready", "False"]],
    [["ready", "True"]], [["Ready", "true"]], [["READY", "TRUE"]], [["Ready ", "True"]],
    [["Ready", "True "]], [["Read", "True"]], [["Ready", "Tru"]], [["Readx", "True"]],
    [["Ready", "Trux"]], [["", ""]], [["Ready", ""]], [["", "True"]], [["True", "Ready"]],
    [["plain:Ready", "True"]], [["plain:Ready", "True"], ["Ready", "False"]],
    [["plain:Ready", "False"], ["Ready", "True"]],
    _PRESSURE + [["Ready", "True"]], _PRESSURE + [["Ready", "Unknown"]],
    [[f"Custom{j}", "False"] for j in range(11)] + [["Ready", "True"]],
    [["X" * 255, "True"], ["Ready", "True"]], [["Ready", "T" * 300]], [["R" * 260, "T" * 260], ["Ready", "True"]],
    [[None, None]], [["Ready", None]], [[None, "True"]],
]

_BASE_CAP = [["cpu", "64"], ["ephemeral-storage", "203070420Ki"], ["hugepages-1Gi", "0"],
             ["hugepages-2Mi", "0"], ["memory", "527951004Ki"], ["pods", "110"]]


def _with(cap_extra: Sequence[Sequence[Any]], conds: Any = "ready") -> Dict[str, Any]:
    cap = [list(p) for p in _BASE_CAP] + [list(p) for p in cap_extra]
    return {"capacity": cap,
            "conditions": (_PRESSURE + [["Ready", "True"]]) if conds == "ready" else conds}


def g2_adversarial_templates() -> List[Dict[str, Any]]:
    t: List[Dict[str, Any]] = []
    keys = pnl.GPU_RESOURCE_KEYS
    for j, q in enumerate(QUANTITY_CASES):                       # every quantity on a rotating key
        t.append(_with([[keys[j % 4], q]]))
        t.append(_with([[keys[(j + 1) % 4], q], [keys[(j + 2) % 4], "2"]], [["Ready", "False"]]))
    for d in KEY_DECOYS:                                         # decoy keys, alone and beside a real one
        t.append(_with([[d, "4"]]))
        t.append(_with([[d, "4"], ["amd.com/gpu", "1"]]))
    for c in CONDITION_CASES:                                    # Ready evaluation
        t.append(_with([["nvidia.com/gpu", "8"]], c))
        t.append(_with([], c))
    multi = [
        [["nvidia.com/gpu", "4"], ["amd.com/gpu", "2"]],
        [["intel.com/gpu", "1"], ["amd.com/gpu", "2"], ["nvidia.com/gpu", "0"]],
        [["nvidia.com/gpu", "8"], ["amd.com/gpu", "8"], ["gpu.intel.com/i915", "8"], ["intel.com/gpu", "8"]],
        [["nvidia.com/gpu", "-1"], ["amd.com/gpu", "2"]], [["nvidia.com/gpu", "-1"], ["amd.com/gpu", "1"]],
        [["nvidia.com/gpu", "-3"], ["amd.com/gpu", "2"]], [["nvidia.com/gpu", "0"], ["amd.com/gpu", "0"]],
        [["nvidia.com/gpu", "1" * 23], ["amd.com/gpu", "-" + "1" * 23]],
        [["nvidia.com/gpu", "1" * 23], ["amd.com/gpu", "-5"]],
        [["nvidia.com/gpu", "-" + "1" * 23], ["amd.com/gpu", "5"]],
        [["nvidia.com/gpu", "9" * 18], ["amd.com/gpu", "9" * 18], ["gpu.intel.com/i915", "9" * 18],
         ["intel.com/gpu", "9" * 18]],
        [["nvidia.com/gpu", "1" + "0" * 18], ["amd.com/gpu", "-" + "9" * 18]],
        [["nvidia.com/gpu", "1k"], ["amd.com/gpu", "2"]], [["nvidia.com/gpu", "1k"], ["amd.com/gpu", "2Ki"]],
        [["nvidia.com/gpu", ""], ["intel.com/gpu", "3"]], [["nvidia.com/gpu", None], ["intel.com/gpu", "3"]],
    ]
    for m in multi:
        t.append(_with(m))
        t.append(_with(m, [["Ready", "Unknown"]]))
    t.append({"capacity": None, "conditions": [["Ready", "True"]]})
    t.append({"capacity": [], "conditions": [["Ready", "True"]]})
    t.append({"status": False})
    t.append({"status": False, "metadata": False})
    t.append({"capacity": [["nvidia.com/gpu", "2"]], "conditions": [["Ready", "True"]], "metadata": False})
    t.append({"capacity": [["nvidia.com/gpu", "2"]], "conditions": None, "spec": False})
    t.append({"capacity": [[f"vendor{j}.example.com/dev", str(j)] for j in range(40)] + [["nvidia.com/gpu", "1"]],
              "conditions": [["Ready", "True"]]})
    t.append({"capacity": [["nvidia.com/gpu", "1"]] + [[f"r{j}", "1"] for j in range(300)],
              "conditions": [["Ready", "True"]]})                       # cap_len > 255 entries
    t.append({"capacity": [["x" * 40000, "1"], ["y" * 30000, "2"], ["amd.com/gpu", "3"]],
              "conditions": [["Ready", "True"]]})                       # cap_len >= 0xFFFF escape
    t.append({"capacity": [["nvidia.com/gpu", " " * 70000 + "6"]], "conditions": [["Ready", "True"]]})
    return t


def g2_templates() -> List[Dict[str, Any]]:
    """G1's 1224 ordinary templates followed by the adversarial ones."""
    return g1_templates() + g2_adversarial_templates()


def g2_tid(seed: int, start: int, n: int, n_adv: Optional[int] = None) -> np.ndarray:
    """Half the nodes are ordinary G1 draws, half are uniform draws from the adversarial list."""
    n_adv = n_adv if n_adv is not None else len(g2_adversarial_templates())
    h = node_hash(seed ^ 0x5EED, start, n)
    adv = (h >> np.uint64(63)).astype(bool)
    pick = ((h >> np.uint64(20)) % np.uint64(n_adv)).astype(np.int32)
    return np.where(adv, 1224 + pick, g1_tid(seed, start, n)).astype(np.int32)


# ----------------------------------------------------------------------------------------------
# Templates -> objects / rows / fleets.
# ----------------------------------------------------------------------------------------------
class _PlainCondition:
    """Has .type/.status but is not a V1NodeCondition (check-gpu-node.py:176 rejects it)."""

    def __init__(self, type, status):
        self.type, self.status = type, status


def build_node(spec: Dict[str, Any], payload: Optional[Dict[str, Any]] = None):
    """Node spec -> `kubernetes.client.V1Node` (the installed client's classes, real or stub)."""
    from kubernetes import client as kc
    p = dict(spec)
    p.update(payload or {})
    status = None
    if p.get("status", True) is not False:
        cap = p.get("capacity")
        conds = p.get("conditions")
        objs = None
        if conds is not None:
            objs = []
            for t, s in conds:
                if isinstance(t, str) and t.startswith("plain:"):
                    objs.append(_PlainCondition(t[6:], s))
                else:
                    objs.append(kc.V1NodeCondition(type=t, status=s))
        status = kc.V1NodeStatus(capacity=None if cap is None else {k: v for k, v in cap}, conditions=objs)
    meta = None
    if p.get("metadata", True) is not False:
        meta = kc.V1ObjectMeta(name=p.get("name"), labels=p.get("labels"))
    sp = None
    if p.get("spec", True) is not False:
        taints = p.get("taints")
        sp = kc.V1NodeSpec(taints=None if taints is None else [kc.V1Taint(**t) for t in taints])
    return kc.V1Node(metadata=meta, spec=sp, status=status)


def template_rows(templates: Sequence[Dict[str, Any]]) -> List[bytes]:
    """PNL-v2 row of every template, through the same packer real nodes go through."""
    return [pnl.node_row(build_node(t)) for t in templates]


def spec_row(spec: Dict[str, Any]) -> bytes:
    """PNL-v2 row of a node spec without building client objects (so callers that only need bytes --
    the benchmark's GPU arm -- do not need a `kubernetes` package).  Same marshalling as
    `pnl.node_row(build_node(spec))`: tests/test_oracle_pnl.py holds the two equal on every template."""
    if spec.get("status", True) is False:
        return pnl.EMPTY_ROW
    cap = spec.get("capacity")
    conds = spec.get("conditions")
    pairs = None
    if cap:                                                  # dict(cap): a repeated key keeps its first position, last value
        d: Dict[Any, Any] = {}
        for k, v in cap:
            d[k] = v
        pairs = list(d.items())
    cl = None
    if conds:                                                # "plain:" entries are not V1NodeCondition objects: left out
        cl = [(t, st) for t, st in conds if not (isinstance(t, str) and t.startswith("plain:"))]
    return pnl.encode_row(pairs, cl)


def spec_rows(templates: Sequence[Dict[str, Any]]) -> List[bytes]:
    return [spec_row(t) for t in templates]


def expand(rows: Sequence[bytes], tid, device=None):
    """(rows u8[T+pad], row_ptr u32-as-int64... ) -> torch tensors (rows: uint8, row_ptr: int32 view of u32).

    Returns (rows_t, row_ptr_t, n_bytes).  `row_ptr_t` is int64 on return-side arithmetic but stored
    as uint32 bit patterns in an int32 tensor (torch has no general uint32 arithmetic)."""
    import torch
    tid_t = torch.as_tensor(tid, dtype=torch.int64, device=device)
    dev = tid_t.device
    n = tid_t.numel()
    lens = torch.tensor([len(r) for r in rows], dtype=torch.int64, device=dev)
    offs = torch.zeros(len(rows) + 1, dtype=torch.int64, device=dev)
    offs[1:] = torch.cumsum(lens, 0)
    blob = torch.frombuffer(bytearray(b"".join(rows)), dtype=torch.uint8).to(dev)
    row_ptr = torch.zeros(n + 1 + pnl.ROWPTR_PAD, dtype=torch.int64, device=dev)
    if n:
        row_ptr[1:n + 1] = torch.cumsum(lens[tid_t], 0)
    total = int(row_ptr[n].item()) if n else 0
    row_ptr[n + 1:] = total
    if total > pnl.MAX_ROWS_BYTES:
        raise ValueError("fleet exceeds the 4 GiB PNL-v2 row limit; generate it in slices")
    out = torch.zeros(total + pnl.ROWS_PAD, dtype=torch.uint8, device=dev)
    chunk = 1 << 20                                     # nodes per pass: bounds the index temporaries
    for lo in range(0, n, chunk):
        hi = min(n, lo + chunk)
        t = tid_t[lo:hi]
        ln = lens[t]
        start = row_ptr[lo:hi]
        b0, b1 = int(row_ptr[lo].item()), int(row_ptr[hi].item())
        node = torch.repeat_interleave(torch.arange(hi - lo, device=dev), ln)
        pos = torch.arange(b0, b1, device=dev)
        out[b0:b1] = blob[offs[t][node] + (pos - start[node])]
    # uint32 bit patterns in an int32 tensor
    rp32 = (row_ptr & 0xFFFFFFFF).to(torch.int64)
    rp32 = torch.where(rp32 >= 2 ** 31, rp32 - 2 ** 32, rp32).to(torch.int32)
    return out, rp32, total


def fleet(kind: str, seed: int, start: int, n: int):
    """(templates, tid) for 'g1' or 'g2'."""
    if kind == "g1":
        return g1_templates(), g1_tid(seed, start, n)
    if kind == "g2":
        return g2_templates(), g2_tid(seed, start, n)
    raise ValueError(kind)


# ----------------------------------------------------------------------------------------------
# Config 5: fleets rendered as a raw /api/v1/nodes document.
# ----------------------------------------------------------------------------------------------
def node_json(spec: Dict[str, Any], payload: Optional[Dict[str, Any]] = None, style: int = 0) -> str:
    """One element of `items` as the API server writes it (compact, Go struct field order).  `style`
    > 0 decorates it with things a conforming parser must see through: whitespace, escaped key and
    value spellings, duplicate keys (last wins), nested values to skip, non-string quantities."""
    import json as _json
    p = dict(spec)
    p.update(payload or {})
    def _u(s):                                                            # every character as \\uXXXX (pairs above the BMP)
        out = []
        for c in s:
            cp = ord(c)
            if cp > 0xFFFF:
                cp -= 0x10000
                out.append(f"\\u{0xD800 + (cp >> 10):04x}\\u{0xDC00 + (cp & 0x3FF):04x}")
            else:
                out.append(f"\\u{cp:04x}")
        return "".join(out)
    esc = _u if style & 2 else (lambda s: _json.dumps(s)[1:-1])
    q = lambda s: '"' + (esc(s) if isinstance(s, str) else "") + '"'      # noqa: E731
    sp = " \n\t " if style & 1 else ""
    parts = []
    if p.get("metadata", True) is not False:
        md = {"name": p.get("name"), "labels": p.get("labels"), "annotations": {"note": "a \"quoted\" {brace} [bracket] \\ path"}}
        parts.append('"metadata":' + sp + _json.dumps(md, ensure_ascii=bool(style & 2)))
    if p.get("spec", True) is not False:
        parts.append('"spec":' + _json.dumps({"taints": p.get("taints")}))
    if p.get("status", True) is not False:
        st = []
        cap = p.get("capacity")
        if style & 4 and cap is not None:                                 # decoys that must lose: an earlier duplicate, a look-alike
            st.append('"capacity":{"nvidia.com/gpu":"64"}')
            st.append('"capacityX":{"amd.com/gpu":"9"}')
        if cap is not None:
            ents = []
            for k, v in cap:
                if isinstance(v, str):
                    ents.append(q(k) + ":" + sp + q(v))
                else:
                    ents.append(q(k) + ":" + _json.dumps(v))
            st.append('"allocatable":{"nvidia.com/gpu":"1000"}')
            st.append('"capacity":' + sp + "{" + ("," + sp).join(ents) + "}")
        conds = p.get("conditions")
        if conds is not None:
            cs = []
            for t, s_ in conds:
                if isinstance(t, str) and t.startswith("plain:"):
                    cs.append(_json.dumps([t[6:], s_]))                  # not an object: no V1NodeCondition comes of it
                    continue
                body = ['"lastHeartbeatTime":"2025-01-01T00:00:00Z"', '"type":' + (q(t) if isinstance(t, str) else "null"),
                        '"status":' + (q(s_) if isinstance(s_, str) else "null"), '"reason":"Kubelet{Ready}","message":"ok, \\"fine\\""']
                if style & 4:
                    body.insert(0, '"status":"True","type":"Ready"')     # overwritten by the later type/status
                cs.append("{" + ",".join(body) + "}")
            st.append('"conditions":' + sp + "[" + ("," + sp).join(cs) + "]")
        st.append('"nodeInfo":{"kubeletVersion":"v1.31.0","nested":{"deep":[1,2,{"x":"}]"}]}}')
        parts.append('"status":' + sp + "{" + ",".join(st) + "}")
    return "{" + ("," + sp).join(parts) + "}"


def nodelist_json(templates: Sequence[Dict[str, Any]], tid, payload_fn=None, style_fn=None) -> bytes:
    """A NodeList document for the fleet drawn from `templates` by `tid`."""
    items = []
    for i, t in enumerate(tid):
        items.append(node_json(templates[int(t)], payload_fn(i, templates[int(t)]) if payload_fn else {"name": f"node-{i:08d}"},
                               style_fn(i) if style_fn else 0))
    head = '{"kind":"NodeList","apiVersion":"v1","metadata":{"resourceVersion":"12345","extra":[{"a":1},{"b":[{}]}]},"items":['
    return (head + ",".join(items) + "]}").encode("utf-8", "surrogatepass")
