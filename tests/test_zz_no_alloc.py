"""gnc_classify_static: the entry point for callers that may not allocate device memory.

Such a caller (an agent running under a device-memory cap it has already used up, or one that
cannot afford an allocation's device-wide synchronisation on its path) still hands in its own input
and output buffers; what it cannot do is let the library cudaMalloc the kernel's scratch.  This
entry point keeps that scratch in a fixed array inside libgnc.so instead.  The test calls it through
the C ABI twice in a row -- a larger fleet, then a smaller, different one, so anything the first call
left in the shared array would show -- and requires every output to equal gnc_classify's, bit for
bit (integer and byte work: no tolerance).  Collected last on purpose; passes alone or after the
rest of the suite.
"""
import ctypes

import numpy as np
import pytest
import torch

from k8s_gpu_node_checker_b200 import engine, native, pnl, synth

pytestmark = pytest.mark.gpu


def _fleet(seed, n):
    tmpl, tid = synth.fleet("g2", seed, 0, n)
    rows = synth.template_rows(tmpl)
    return pnl.from_rows([rows[t] for t in tid])


def _host(out: engine.DeviceResult, n: int):
    counts = out.counts.cpu().numpy()
    g, d = int(counts[0]), min(int(counts[2]), out.defer_cap)
    defer = out.defer.cpu().numpy().view(engine._DEFER_DT)[:d]
    return {"counts": counts, "flags": out.flags.cpu().numpy()[:n], "gpu_idx": out.gpu_idx.cpu().numpy()[:g],
            "gpu_total": out.gpu_total.cpu().numpy()[:g], "gpu_vals": out.gpu_vals.cpu().numpy()[:g],
            "defer": np.sort(defer, order="node")}


def _classify_static(lib, batch: engine.DeviceBatch, out: engine.DeviceResult, max_digits=4300):
    b = native.Batch(batch.rows.data_ptr(), batch.row_ptr.data_ptr(), batch.rows_bytes, batch.n_nodes, batch.node_base)
    r = out.c_struct()
    err = ctypes.create_string_buffer(256)
    stream = torch.cuda.current_stream().cuda_stream
    rc = lib.gnc_classify_static(0, ctypes.byref(b), ctypes.byref(r), max_digits, ctypes.c_void_p(stream), err, len(err))
    return rc, err.value.decode()


def test_static_scratch_entry_point_matches_gnc_classify():
    lib = native.lib()
    dev = torch.device("cuda", 0)
    clf = engine.Classifier(0)
    try:
        for seed, n in ((31, 60000), (32, 7001)):
            p = _fleet(seed, n)
            batch = engine.DeviceBatch.from_packed(p, dev, node_base=17)
            want = engine.DeviceResult.alloc(n, dev, defer_cap=1 << 14)
            clf.classify(batch, want)
            got = engine.DeviceResult.alloc(n, dev, defer_cap=1 << 14)
            got.flags.fill_(0xEE)
            rc, msg = _classify_static(lib, batch, got)
            assert rc == 0, msg
            torch.cuda.synchronize()
            a, b = _host(got, n), _host(want, n)
            assert int(a["counts"][0]) > 0 and int(a["counts"][2]) > 0       # GPU nodes and bignum nodes both present
            for key in ("counts", "flags", "gpu_idx", "gpu_total", "gpu_vals"):
                np.testing.assert_array_equal(a[key], b[key], err_msg=f"{key}, fleet seed {seed}")
            np.testing.assert_array_equal(a["defer"], b["defer"])
    finally:
        clf.close()


def test_static_scratch_entry_point_refuses_oversize_shards():
    lib = native.lib()
    dev = torch.device("cuda", 0)
    rows = torch.zeros(4096, dtype=torch.uint8, device=dev)
    rp = torch.zeros(4096, dtype=torch.int32, device=dev)
    out = engine.DeviceResult.alloc(16, dev)
    n = native.STATIC_MAX_TILES * 256 + 1                       # one node more than the array's descriptors cover (256-node tiles)
    big = engine.DeviceBatch(rows, rp, n, n * 200)
    rc, msg = _classify_static(lib, big, out)
    assert rc == -1 and "static scratch holds 65536" in msg, msg
    rc, msg = _classify_static(lib, engine.DeviceBatch(rows[1:], rp, 8, 32), out)
    assert rc == -1 and "aligned" in msg
