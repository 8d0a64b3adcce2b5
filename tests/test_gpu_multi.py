"""Two ranks, two GPUs, through gnc_gather (NCCL): the combined result equals the single-GPU one.
Skipped on boxes with fewer than two GPUs (the round-end `-m gpu` run has one)."""
import os
import socket

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, out_dir):
    import torch.distributed as dist
    from oracle import ensure_kubernetes_importable
    ensure_kubernetes_importable()
    from k8s_gpu_node_checker_b200 import engine, pnl, sharded, synth
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)        # only to ship the NCCL id
    box = [engine.Classifier.comm_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    clf = engine.Classifier(rank, rank, world, box[0])
    tmpl, _ = synth.fleet("g2", 77, 0, 1)
    rows = synth.template_rows(tmpl)
    lo, hi = sharded.shard_range(n, rank, world)
    _, tid = synth.fleet("g2", 77, lo, hi - lo)
    p = pnl.from_rows([rows[t] for t in tid])
    dev = torch.device("cuda", rank)
    local = clf.classify(engine.DeviceBatch.from_packed(p, dev, node_base=lo),
                         engine.DeviceResult.alloc(hi - lo, dev, defer_cap=1 << 15))
    fleet = engine.DeviceResult.alloc(n, dev, defer_cap=1)
    lean = engine.DeviceResult.alloc(n, dev, defer_cap=1, records=False)   # indices + counts only: the sync-free path
    for _ in range(2):                                                  # twice: the gather is repeatable
        clf.gather(local, fleet)
        clf.gather(local, lean)
    torch.cuda.synchronize()
    assert torch.equal(lean.counts, fleet.counts)
    assert torch.equal(lean.gpu_idx[:int(lean.counts[0])], fleet.gpu_idx[:int(fleet.counts[0])])
    # the fused gather: classify itself stores the indices into every rank over NVLink; run it for several
    # epochs so that both buffer parities are reused and the acknowledgements are exercised
    clf.fleet_enable(n)
    fused = engine.DeviceResult.alloc(n, dev, defer_cap=1, records=False)
    for _ in range(5):
        fused.gpu_idx.fill_(-1)
        clf.classify(engine.DeviceBatch.from_packed(p, dev, node_base=lo), local)
        clf.gather(local, fused)
        torch.cuda.synchronize()
        assert torch.equal(fused.counts, fleet.counts), (fused.counts, fleet.counts)
        assert torch.equal(fused.gpu_idx[:int(fused.counts[0])], fleet.gpu_idx[:int(fleet.counts[0])])
    counts = fleet.counts.cpu().numpy()
    g = int(counts[0])
    np.save(os.path.join(out_dir, f"c{rank}.npy"), counts)
    np.save(os.path.join(out_dir, f"i{rank}.npy"), fleet.gpu_idx.cpu().numpy().view(np.uint32)[:g])
    np.save(os.path.join(out_dir, f"t{rank}.npy"), fleet.gpu_total.cpu().numpy()[:g])
    np.save(os.path.join(out_dir, f"v{rank}.npy"), fleet.gpu_vals.cpu().numpy()[:g])
    clf.close()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("variant", [None, "t224", "t128", "t32"])      # the tile records of the fused gather depend on the tile size
def test_two_rank_gather_equals_single_gpu(tmp_path, variant, monkeypatch):
    import torch.multiprocessing as mp
    from k8s_gpu_node_checker_b200 import engine, pnl, synth
    n, world = 300_001, 2
    if variant:
        monkeypatch.setenv("GNC_VARIANT", variant)                      # read by gnc_create in the spawned ranks
    mp.spawn(_worker, args=(world, _free_port(), n, str(tmp_path)), nprocs=world, join=True)
    monkeypatch.delenv("GNC_VARIANT", raising=False)
    tmpl, tid = synth.fleet("g2", 77, 0, n)
    rows = synth.template_rows(tmpl)
    p = pnl.from_rows([rows[t] for t in tid])
    clf = engine.Classifier(0)
    out = clf.classify(engine.DeviceBatch.from_packed(p, clf.device), engine.DeviceResult.alloc(n, clf.device, defer_cap=1 << 16))
    torch.cuda.synchronize()
    counts = out.counts.cpu().numpy()
    g = int(counts[0])
    for r in range(world):
        np.testing.assert_array_equal(np.load(tmp_path / f"c{r}.npy")[:13], counts[:13])
        np.testing.assert_array_equal(np.load(tmp_path / f"i{r}.npy"), out.gpu_idx.cpu().numpy().view(np.uint32)[:g])
        np.testing.assert_array_equal(np.load(tmp_path / f"t{r}.npy"), out.gpu_total.cpu().numpy()[:g])
        np.testing.assert_array_equal(np.load(tmp_path / f"v{r}.npy"), out.gpu_vals.cpu().numpy()[:g])
    clf.close()
