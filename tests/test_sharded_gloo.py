"""World-size-2 (and 3) runs of the shard/combine logic over gloo on the CPU.

The per-shard classification is stood in for by the C oracle (this is test code; the product's ranks
run the CUDA kernel) -- what is under test is the host-side N>1 path: shard ranges, node_base offsets,
the counts reduction (128-bit key totals) and the rank-ordered concatenation of compacted indices."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from k8s_gpu_node_checker_b200 import pnl, sharded, synth
from oracle import pnl_oracle as po


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n, kind, seed, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        tmpl, _ = synth.fleet(kind, seed, 0, 1)
        rows = synth.template_rows(tmpl)
        lo, hi = sharded.shard_range(n, rank, world)
        _, tid = synth.fleet(kind, seed, lo, hi - lo)               # each rank regenerates only its slice
        p = pnl.from_rows([rows[t] for t in tid])
        local = po.classify(p.rows, p.row_ptr, hi - lo, node_base=lo)
        res = sharded.gather_fleet(torch.from_numpy(local["counts"]),
                                   torch.from_numpy(local["gpu_idx"].astype(np.int64)))
        np.save(os.path.join(out_dir, f"idx{rank}.npy"), res["gpu_idx"].numpy())
        np.save(os.path.join(out_dir, f"counts{rank}.npy"), res["counts"].numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n,kind", [(2, 5001, "g1"), (2, 4000, "g2"), (3, 1000, "g2")])
def test_sharded_equals_single(tmp_path, world, n, kind):
    seed = 5
    port = _free_port()
    mp.spawn(_worker, args=(world, port, n, kind, seed, str(tmp_path)), nprocs=world, join=True)
    tmpl, tid = synth.fleet(kind, seed, 0, n)
    rows = synth.template_rows(tmpl)
    p = pnl.from_rows([rows[t] for t in tid])
    whole = po.classify(p.rows, p.row_ptr, n)
    for r in range(world):
        idx = np.load(tmp_path / f"idx{r}.npy")
        counts = np.load(tmp_path / f"counts{r}.npy")
        np.testing.assert_array_equal(idx, whole["gpu_idx"].astype(np.int64))      # one ordered list, on every rank
        np.testing.assert_array_equal(counts[:13], whole["counts"][:13])
        assert sharded.fleet_exit_code(counts) == (0 if whole["counts"][1] else 3 if whole["counts"][0] else 2)


def test_shard_ranges_cover_the_fleet():
    for n in (0, 1, 7, 10_000_000):
        for w in (1, 2, 3, 4, 8):
            r = [sharded.shard_range(n, k, w) for k in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n and all(a[1] == b[0] for a, b in zip(r, r[1:]))


def test_counts_reduction_is_exact_beyond_int64():
    a = np.zeros(16, np.int64)
    b = np.zeros(16, np.int64)
    a[4], a[8] = -1, 0            # lo = 2^64-1
    b[4], b[8] = 5, 0
    s = sharded.reduce_counts(np.stack([a, b]))
    assert (int(s[8]) << 64) + (int(s[4]) & ((1 << 64) - 1)) == (1 << 64) - 1 + 5
