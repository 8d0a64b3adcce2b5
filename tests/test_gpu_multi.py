"""W ranks on W GPUs: the combined result of every gather form (NCCL with records, NCCL indices + counts,
fused NVLink gather) equals the single-GPU result -- the reference's one ordered list
(check-gpu-node.py:220-225) and the sums one_shot derives (:275-276).  Parametrised over the world sizes
the box can hold (2, 4, 8); skipped where there are fewer GPUs (the round-end `-m gpu` box has one;
`bench.py --gpus N` carries its own check of the same thing, see its "verified" key)."""
import os
import socket

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

N_GPUS = torch.cuda.device_count() if torch.cuda.is_available() else 0


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _shard(rank, world, n, seed=77):
    from k8s_gpu_node_checker_b200 import pnl, sharded, synth
    tmpl, _ = synth.fleet("g2", seed, 0, 1)
    rows = synth.template_rows(tmpl)
    lo, hi = sharded.shard_range(n, rank, world)
    _, tid = synth.fleet("g2", seed, lo, hi - lo)
    return pnl.from_rows([rows[t] for t in tid]), lo, hi


def _worker(rank, world, port, n, out_dir, short_by):
    import torch.distributed as dist
    from k8s_gpu_node_checker_b200 import engine, native
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)        # only to ship the NCCL id
    box = [engine.Classifier.comm_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    clf = engine.Classifier(rank, rank, world, box[0])
    p, lo, hi = _shard(rank, world, n)
    dev = torch.device("cuda", rank)
    batch = engine.DeviceBatch.from_packed(p, dev, node_base=lo)
    local = clf.classify(batch, engine.DeviceResult.alloc(hi - lo, dev, defer_cap=1 << 15))
    fleet = engine.DeviceResult.alloc(n, dev, defer_cap=1)
    lean = engine.DeviceResult.alloc(n, dev, defer_cap=1, records=False)   # indices + counts only: the sync-free path
    for _ in range(2):                                                  # twice: the gather is repeatable
        clf.gather(local, fleet)
        clf.gather(local, lean)
    torch.cuda.synchronize()
    g = int(fleet.counts[0])
    assert torch.equal(lean.counts, fleet.counts)
    assert torch.equal(lean.gpu_idx[:g], fleet.gpu_idx[:g])
    assert int(fleet.counts[13]) == 0 and int(fleet.counts[14]) == 0

    # a gather target that is too small: the two device-side forms clip and say by how much in counts[14];
    # the records form refuses on EVERY rank (the decision is taken from all-gathered numbers, so no rank
    # is left waiting in a collective the others skipped)
    if short_by and g > short_by:
        small = engine.DeviceResult.alloc(n, dev, out_cap=g - short_by, defer_cap=1, records=False)
        clf.gather(local, small)
        torch.cuda.synchronize()
        assert int(small.counts[14]) == short_by and int(small.counts[0]) == g
        assert torch.equal(small.gpu_idx[:g - short_by], fleet.gpu_idx[:g - short_by])
        # only rank 0's table is short: every rank must still agree on the outcome
        small_rec = engine.DeviceResult.alloc(n, dev, out_cap=(g - short_by) if rank == 0 else n, defer_cap=1)
        with pytest.raises(native.GncError) as ei:
            clf.gather(local, small_rec)
        assert ei.value.code == -6
        clf.gather(local, fleet)                                        # and the communicator is still in step
        torch.cuda.synchronize()

    # the fused gather: classify itself stores ballots, prefixes and counts into every rank over NVLink; run it
    # for several epochs so that both buffer parities are reused and the acknowledgements are exercised
    clf.fleet_enable(n)
    fused = engine.DeviceResult.alloc(n, dev, defer_cap=1, records=False)
    for _ in range(5):
        fused.gpu_idx.fill_(-1)
        clf.classify(batch, local)
        clf.gather(local, fused)
        torch.cuda.synchronize()
        assert torch.equal(fused.counts, fleet.counts), (fused.counts, fleet.counts)
        assert torch.equal(fused.gpu_idx[:g], fleet.gpu_idx[:g])
    if short_by and g > short_by:
        small = engine.DeviceResult.alloc(n, dev, out_cap=g - short_by, defer_cap=1, records=False)
        clf.classify(batch, local)
        clf.gather(local, small)
        torch.cuda.synchronize()
        assert int(small.counts[14]) == short_by and int(small.counts[0]) == g
        assert torch.equal(small.gpu_idx[:g - short_by], fleet.gpu_idx[:g - short_by])
    # the same with the target named before the launch: classify rebuilds the list itself, gather has nothing left to do
    onek = engine.DeviceResult.alloc(n, dev, defer_cap=1, records=False)
    clf.fleet_target(onek)
    for _ in range(4):
        onek.gpu_idx.fill_(-1)
        onek.counts.fill_(-1)
        before = clf.launches
        clf.classify(batch, local)
        clf.gather(local, onek)
        torch.cuda.synchronize()
        assert clf.launches - before == (1 if hi > lo else 2)           # one launch did both (an empty shard: its notice, then the rebuild)
        assert torch.equal(onek.counts, fleet.counts), (onek.counts, fleet.counts)
        assert torch.equal(onek.gpu_idx[:g], fleet.gpu_idx[:g])
    clf.gather(local, fused)                                            # another target: the rebuild runs as a launch of its own
    torch.cuda.synchronize()
    assert torch.equal(fused.gpu_idx[:g], fleet.gpu_idx[:g])
    if short_by and g > short_by:
        small = engine.DeviceResult.alloc(n, dev, out_cap=g - short_by, defer_cap=1, records=False)
        clf.fleet_target(small)
        clf.classify(batch, local)
        clf.gather(local, small)
        torch.cuda.synchronize()
        assert int(small.counts[14]) == short_by and int(small.counts[0]) == g
        assert torch.equal(small.gpu_idx[:g - short_by], fleet.gpu_idx[:g - short_by])
    clf.fleet_target(None)
    counts = fleet.counts.cpu().numpy()
    np.save(os.path.join(out_dir, f"c{rank}.npy"), counts)
    np.save(os.path.join(out_dir, f"i{rank}.npy"), fleet.gpu_idx.cpu().numpy().view(np.uint32)[:g])
    np.save(os.path.join(out_dir, f"t{rank}.npy"), fleet.gpu_total.cpu().numpy()[:g])
    np.save(os.path.join(out_dir, f"v{rank}.npy"), fleet.gpu_vals.cpu().numpy()[:g])
    clf.close()
    dist.destroy_process_group()


def _single_gpu(n):
    from k8s_gpu_node_checker_b200 import engine
    p, _, _ = _shard(0, 1, n)
    clf = engine.Classifier(0)
    out = clf.classify(engine.DeviceBatch.from_packed(p, clf.device), engine.DeviceResult.alloc(n, clf.device, defer_cap=1 << 16))
    torch.cuda.synchronize()
    counts = out.counts.cpu().numpy()
    g = int(counts[0])
    res = (counts, out.gpu_idx.cpu().numpy().view(np.uint32)[:g], out.gpu_total.cpu().numpy()[:g], out.gpu_vals.cpu().numpy()[:g])
    clf.close()
    return res


def _check(tmp_path, world, n):
    counts, idx, tot, vals = _single_gpu(n)
    for r in range(world):
        np.testing.assert_array_equal(np.load(tmp_path / f"c{r}.npy")[:13], counts[:13])
        np.testing.assert_array_equal(np.load(tmp_path / f"i{r}.npy"), idx)
        np.testing.assert_array_equal(np.load(tmp_path / f"t{r}.npy"), tot)
        np.testing.assert_array_equal(np.load(tmp_path / f"v{r}.npy"), vals)


# the tile records of the fused gather depend on the tile size (variant) and on the world size (slot stride, flag and
# acknowledgement arrays, peer stores per ballot word)
CASES = [(2, None), (2, "t224"), (2, "t128g2"), (2, "t32"), (4, None), (4, "t64g2"), (8, None), (8, "t192g2")]


@pytest.mark.parametrize("world,variant", CASES)
def test_gather_equals_single_gpu(tmp_path, world, variant, monkeypatch):
    if N_GPUS < world:
        pytest.skip(f"needs {world} GPUs, box has {N_GPUS}")
    import torch.multiprocessing as mp
    n = 300_001
    if variant:
        monkeypatch.setenv("GNC_VARIANT", variant)                      # read by gnc_create in the spawned ranks
    mp.spawn(_worker, args=(world, _free_port(), n, str(tmp_path), 5), nprocs=world, join=True)
    monkeypatch.delenv("GNC_VARIANT", raising=False)
    _check(tmp_path, world, n)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_ranks_with_empty_shards(tmp_path, world):
    """Fewer nodes than ranks: some shards are empty; they still take part in every form of the gather."""
    if N_GPUS < world:
        pytest.skip(f"needs {world} GPUs, box has {N_GPUS}")
    import torch.multiprocessing as mp
    n = world - 1
    mp.spawn(_worker, args=(world, _free_port(), n, str(tmp_path), 0), nprocs=world, join=True)
    _check(tmp_path, world, n)
