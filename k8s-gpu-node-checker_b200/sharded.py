"""Sharded fleets: which nodes a rank owns, and how per-rank results combine.

Rank r of W owns the contiguous range [r*N/W, (r+1)*N/W).  Each rank classifies and compacts its own
shard (indices already carry the shard's `node_base`); combining needs exactly two collectives:

  * the counts blocks of all ranks (their exact sum is the all-reduce of the counts; the per-rank
    GPU-node counts are the gather offsets), and
  * an all-gather-v of the compacted GPU-node indices into rank-ordered slices, which reproduces the
    reference's single ordered list (check-gpu-node.py:220-225).

On GPUs the library does this itself over NCCL (`gnc_gather`).  This module is the same logic over a
`torch.distributed` process group, used for the host-side tests (gloo, CPU) and by callers that
already hold a process group.
"""
from __future__ import annotations

from typing import Dict, List, Tuple

import numpy as np
import torch
import torch.distributed as dist

from .native import C_GPU_NODES, C_KEY_HI, C_KEY_LO, N_COUNTS


def shard_range(n_nodes: int, rank: int, world: int) -> Tuple[int, int]:
    """[lo, hi) of the nodes rank owns; ranges are contiguous, ordered and cover the fleet."""
    return rank * n_nodes // world, (rank + 1) * n_nodes // world


def reduce_counts(blocks: np.ndarray) -> np.ndarray:
    """Exact sum of per-rank counts blocks [W, 16]: plain slots add, key totals add as 128-bit pairs."""
    out = blocks.sum(axis=0).astype(np.int64)
    for k in range(4):
        t = sum((int(b[C_KEY_HI + k]) << 64) + (int(b[C_KEY_LO + k]) & ((1 << 64) - 1)) for b in blocks)
        out[C_KEY_LO + k] = np.array([t & ((1 << 64) - 1)], dtype=np.uint64).view(np.int64)[0]
        out[C_KEY_HI + k] = t >> 64
    return out


def gather_fleet(counts: torch.Tensor, gpu_idx: torch.Tensor, group=None) -> Dict[str, torch.Tensor]:
    """Combine this rank's (counts[16], gpu_idx[>= counts[0]]) with every other rank's.

    Returns {"counts": fleet counts, "gpu_idx": all GPU-node indices in fleet order, "per_rank": [W] counts}.
    Tensors live wherever the inputs live (CPU for gloo, CUDA for nccl)."""
    world = dist.get_world_size(group)
    dev = counts.device
    blocks = [torch.empty(N_COUNTS, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(blocks, counts.contiguous(), group=group)
    host = torch.stack(blocks).cpu().numpy()
    per_rank = [int(b[C_GPU_NODES]) for b in host]
    width = max(per_rank + [1])                                   # all_gather wants equal shapes: pad, then trim
    mine = torch.zeros(width, dtype=gpu_idx.dtype, device=dev)
    mine[:per_rank[dist.get_rank(group)]] = gpu_idx[:per_rank[dist.get_rank(group)]]
    parts = [torch.empty(width, dtype=gpu_idx.dtype, device=dev) for _ in range(world)]
    dist.all_gather(parts, mine, group=group)
    fleet_idx = torch.cat([p[:c] for p, c in zip(parts, per_rank)])        # rank order == fleet order
    return {"counts": torch.from_numpy(reduce_counts(host)), "gpu_idx": fleet_idx, "per_rank": per_rank}


def fleet_exit_code(counts) -> int:
    """check-gpu-node.py:289-293 from the reduced counts (before any bignum nodes are folded in)."""
    if int(counts[1]):
        return 0
    return 3 if int(counts[0]) else 2
