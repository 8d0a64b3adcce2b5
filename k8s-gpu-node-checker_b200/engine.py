"""Host side of the classification path: device buffers, the libgnc calls, the bignum escape.

PyTorch is used for device memory, streams and (in `sharded`) process groups only; every byte of
classification work happens in libgnc.so's CUDA kernels.  There is no CPU path.
"""
from __future__ import annotations

import ctypes
import sys
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from . import native, pnl
from .native import (C_DEFERRED, C_DEFER_DROPPED, C_GPU_NODES, C_KEY_HI, C_KEY_LO, C_OUT_DROPPED, C_READY_NODES,
                     F_GPU, F_KEY0, F_READY, N_COUNTS)

KEYS = pnl.GPU_RESOURCE_KEYS
_DEFER_DT = np.dtype([("node", "<u4"), ("off", "<u4", (4,)), ("len", "<u4", (4,))])


@dataclass
class DeviceBatch:
    """A PNL-v2 shard resident in HBM."""
    rows: torch.Tensor          # uint8, padded
    row_ptr: torch.Tensor       # int32 holding uint32 bit patterns, padded
    n_nodes: int
    rows_bytes: int
    node_base: int = 0

    @staticmethod
    def from_packed(p: pnl.PackedNodeList, device, node_base: int = 0, pin: bool = False) -> "DeviceBatch":
        rows = torch.from_numpy(p.rows)
        rp = torch.from_numpy(p.row_ptr.view(np.int32))
        if pin:
            rows, rp = rows.pin_memory(), rp.pin_memory()
        return DeviceBatch(rows.to(device, non_blocking=pin), rp.to(device, non_blocking=pin), p.n_nodes, p.n_bytes,
                           node_base)


@dataclass
class DeviceResult:
    flags: torch.Tensor
    gpu_idx: torch.Tensor       # int32 holding uint32
    gpu_total: torch.Tensor
    gpu_vals: torch.Tensor      # [cap, 4]
    defer: torch.Tensor         # uint8 view of gnc_defer_rec[defer_cap]
    counts: torch.Tensor
    out_cap: int
    defer_cap: int

    @staticmethod
    def alloc(n_nodes: int, device, out_cap: Optional[int] = None, defer_cap: int = 4096,
              records: bool = True) -> "DeviceResult":
        """`records=False` leaves gpu_total / gpu_vals out (a gather target that only wants the ordered indices)."""
        cap = max(1, n_nodes if out_cap is None else out_cap)
        return DeviceResult(torch.empty(max(1, n_nodes), dtype=torch.uint8, device=device),
                            torch.empty(cap, dtype=torch.int32, device=device),
                            torch.empty(cap if records else 0, dtype=torch.int64, device=device),
                            torch.empty((cap if records else 0, 4), dtype=torch.int64, device=device),
                            torch.empty(max(1, defer_cap) * _DEFER_DT.itemsize, dtype=torch.uint8, device=device),
                            torch.zeros(N_COUNTS, dtype=torch.int64, device=device), cap, defer_cap)

    def c_struct(self) -> native.Result:
        return native.Result(self.flags.data_ptr(), self.gpu_idx.data_ptr(),
                             self.gpu_total.data_ptr() if self.gpu_total.numel() else None,
                             self.gpu_vals.data_ptr() if self.gpu_vals.numel() else None,
                             self.defer.data_ptr(), self.counts.data_ptr(), self.out_cap, self.defer_cap)


@dataclass
class FleetResult:
    """Host-side result for a whole fleet, bignum nodes already folded in."""
    n_nodes: int
    flags: np.ndarray                         # u8[n]
    gpu_idx: np.ndarray                       # i64[G], ascending fleet indices of GPU nodes
    gpu_total: List[int]                      # python ints (arbitrary precision)
    gpu_vals: np.ndarray                      # i64[G,4]; rows of bignum nodes are zero, see `wide`
    key_totals: List[int]                     # exact per-key totals over GPU nodes
    wide: Dict[int, Dict[str, int]] = field(default_factory=dict)   # node -> gpu_breakdown for bignum nodes
    counts: Optional[np.ndarray] = None
    node_base: int = 0                        # fleet index of flags[0]

    @property
    def total_nodes(self) -> int:             # check-gpu-node.py:275
        return len(self.gpu_idx)

    def is_ready(self, pos: int) -> bool:
        return bool(self.flags[int(self.gpu_idx[pos]) - self.node_base] & F_READY)

    @property
    def ready_nodes(self) -> int:             # check-gpu-node.py:276
        return int(((self.flags[self.gpu_idx - self.node_base] & F_READY) != 0).sum()) if len(self.gpu_idx) else 0

    @property
    def exit_code(self) -> int:               # check-gpu-node.py:289-293
        if self.ready_nodes:
            return 0
        return 3 if self.total_nodes else 2

    def breakdown(self, pos: int) -> Dict[str, int]:
        """gpu_breakdown of the pos-th GPU node, keys in GPU_RESOURCE_KEYS order (check-gpu-node.py:186-195)."""
        node = int(self.gpu_idx[pos])
        if node in self.wide:
            return dict(self.wide[node])
        f = int(self.flags[node - self.node_base])
        return {k: int(self.gpu_vals[pos, j]) for j, k in enumerate(KEYS) if f & (F_KEY0 << j)}


def variant_for(rows_bytes: int, n_nodes: int) -> str:
    """Name of the kernel variant gnc_classify picks for a shard of this shape (GNC_VARIANT unset)."""
    return native.lib().gnc_variant_for(int(rows_bytes), int(n_nodes)).decode()


def key_total(counts: np.ndarray, k: int) -> int:
    return (int(counts[C_KEY_HI + k]) << 64) + (int(counts[C_KEY_LO + k]) & ((1 << 64) - 1))


class Classifier:
    """One libgnc context on one GPU."""

    def __init__(self, device: int = 0, rank: int = 0, world: int = 1, nccl_id: Optional[bytes] = None):
        self.lib = native.lib()
        self.device = torch.device("cuda", device)
        self._h = ctypes.c_void_p()
        idbuf = ctypes.create_string_buffer(nccl_id, len(nccl_id)) if nccl_id else None
        rc = self.lib.gnc_create(ctypes.byref(self._h), device, rank, world, idbuf, len(nccl_id) if nccl_id else 0)
        if rc:
            msg = self.lib.gnc_last_error(self._h).decode() if self._h else "allocation failed"
            if self._h:
                self.lib.gnc_destroy(self._h)
                self._h = ctypes.c_void_p()
            raise native.GncError(rc, msg)
        self.rank, self.world = rank, world
        self._check(self.lib.gnc_set_max_str_digits(self._h, sys.get_int_max_str_digits()))

    @staticmethod
    def comm_id() -> bytes:
        buf = ctypes.create_string_buffer(128)
        n = ctypes.c_size_t()
        rc = native.lib().gnc_comm_id(buf, 128, ctypes.byref(n))
        if rc:
            raise native.GncError(rc, "gnc_comm_id failed (is libnccl.so.2 loadable?)")
        return buf.raw[:n.value]

    def _check(self, rc: int):
        if rc:
            raise native.GncError(rc, self.lib.gnc_last_error(self._h).decode())

    def close(self):
        if self._h:
            self.lib.gnc_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def launches(self) -> int:
        return int(self.lib.gnc_launch_count(self._h))

    def reserve(self, max_nodes: int):
        self._check(self.lib.gnc_reserve(self._h, max_nodes))

    # ---- device-resident path ---------------------------------------------------------------
    def classify(self, batch: DeviceBatch, out: Optional[DeviceResult] = None,
                 stream: Optional[torch.cuda.Stream] = None) -> DeviceResult:
        """Enqueue the classification of a device-resident shard; returns without waiting."""
        if out is None:
            out = DeviceResult.alloc(batch.n_nodes, self.device)
        s = stream or torch.cuda.current_stream(self.device)
        b = native.Batch(batch.rows.data_ptr(), batch.row_ptr.data_ptr(), batch.rows_bytes, batch.n_nodes, batch.node_base)
        r = out.c_struct()
        self._check(self.lib.gnc_classify(self._h, ctypes.byref(b), ctypes.byref(r), ctypes.c_void_p(s.cuda_stream)))
        return out

    def fleet_enable(self, fleet_nodes_cap: int, stream: Optional[torch.cuda.Stream] = None):
        """Collective: switch to the fused gather (classify stores indices into every rank over NVLink)."""
        s = stream or torch.cuda.current_stream(self.device)
        self._check(self.lib.gnc_fleet_enable(self._h, fleet_nodes_cap, ctypes.c_void_p(s.cuda_stream)))

    def fleet_target(self, fleet: Optional[DeviceResult]):
        """Name the gather target before the classify launches: classify then does the gather in the same launch
        (every rank must do likewise).  `None` clears it."""
        if fleet is None:
            self._check(self.lib.gnc_fleet_target(self._h, None))
        else:
            self._fleet_target = fleet                                  # keep the buffers alive
            r = fleet.c_struct()
            self._check(self.lib.gnc_fleet_target(self._h, ctypes.byref(r)))

    def fleet_timing(self) -> Dict[str, float]:
        """Mean flag wait and mean / longest rebuild time per CTA (us) since the last call (in-kernel timestamps)."""
        t = (ctypes.c_uint64 * 8)()
        self._check(self.lib.gnc_fleet_timing(self._h, t))
        k = max(1, int(t[3]))
        return {"wait_us": t[0] / k / 1e3, "rebuild_us": t[1] / k / 1e3, "rebuild_us_max": t[2] / 1e3, "ctas": int(t[3]),
                "classify_us": t[4] / k / 1e3, "rebuild_header_us": t[5] / k / 1e3, "rebuild_groups_us": t[6] / k / 1e3}

    def gather(self, local: DeviceResult, fleet: DeviceResult, stream: Optional[torch.cuda.Stream] = None) -> DeviceResult:
        s = stream or torch.cuda.current_stream(self.device)
        a, b = local.c_struct(), fleet.c_struct()
        self._check(self.lib.gnc_gather(self._h, ctypes.byref(a), ctypes.byref(b), ctypes.c_void_p(s.cuda_stream)))
        return fleet

    # ---- raw JSON ingestion -------------------------------------------------------------------
    def json_to_pnl(self, doc: torch.Tensor, nodes_cap: int, rows_cap: Optional[int] = None,
                    stream: Optional[torch.cuda.Stream] = None) -> Tuple[DeviceBatch, torch.Tensor]:
        """Raw /api/v1/nodes bytes (uint8 tensor on this GPU) -> (PNL-v2 shard in HBM, info[8]).

        Synchronises once to read the node count and row bytes; raises on any status bit."""
        s = stream or torch.cuda.current_stream(self.device)
        rows_cap = int(rows_cap if rows_cap is not None else doc.numel() + 64)
        rows = torch.empty(rows_cap, dtype=torch.uint8, device=self.device)
        row_ptr = torch.empty(nodes_cap + 8, dtype=torch.int32, device=self.device)
        info = torch.zeros(8, dtype=torch.int64, device=self.device)
        out = native.JsonOut(rows.data_ptr(), row_ptr.data_ptr(), info.data_ptr(), rows_cap, nodes_cap)
        self._check(self.lib.gnc_json_to_pnl(self._h, ctypes.c_void_p(doc.data_ptr()), doc.numel(), ctypes.byref(out),
                                             ctypes.c_void_p(s.cuda_stream)))
        h = info.cpu()
        status = int(h[2])
        if status:
            names = {1: 'no root "items" array', 2: "nodes_cap too small", 4: "rows_cap too small", 8: "unbalanced JSON",
                     16: "too many root-level arrays"}
            raise native.GncError(-1, "gnc_json_to_pnl: " + ", ".join(v for k, v in names.items() if status & k))
        return DeviceBatch(rows, row_ptr, int(h[0]), int(h[1])), h

    def json_to_pnl_into(self, doc: torch.Tensor, nodes_cap: int, rows: torch.Tensor, row_ptr: torch.Tensor,
                         info: torch.Tensor, stream: Optional[torch.cuda.Stream] = None) -> None:
        """The same scan into caller-held buffers, enqueued without waiting (gnc_json_to_pnl as it is):
        `rows` uint8[rows_cap], `row_ptr` int32[nodes_cap + 8], `info` int64[8] -- read info[0..2] after a sync."""
        s = stream or torch.cuda.current_stream(self.device)
        out = native.JsonOut(rows.data_ptr(), row_ptr.data_ptr(), info.data_ptr(), rows.numel(), nodes_cap)
        self._check(self.lib.gnc_json_to_pnl(self._h, ctypes.c_void_p(doc.data_ptr()), doc.numel(), ctypes.byref(out),
                                             ctypes.c_void_p(s.cuda_stream)))

    # ---- host-buffer path (what a binding without device memory of its own calls) ------------
    def classify_host(self, p: pnl.PackedNodeList, node_base: int = 0, defer_cap: int = 4096,
                      out: Optional[Dict[str, np.ndarray]] = None) -> Dict[str, np.ndarray]:
        n = p.n_nodes
        if out is None:
            out = host_buffers(n, defer_cap)
        b = native.Batch(p.rows.ctypes.data, p.row_ptr.ctypes.data, p.n_bytes, n, node_base)
        r = native.Result(out["flags"].ctypes.data, out["gpu_idx"].ctypes.data, out["gpu_total"].ctypes.data,
                          out["gpu_vals"].ctypes.data, out["defer"].ctypes.data, out["counts"].ctypes.data,
                          len(out["gpu_idx"]), len(out["defer"]))
        self._check(self.lib.gnc_classify_host(self._h, ctypes.byref(b), ctypes.byref(r)))
        return out


def host_buffers(n: int, defer_cap: int = 4096, pinned: bool = False) -> Dict[str, np.ndarray]:
    cap = max(1, n)
    shapes = {"flags": (cap, np.uint8), "gpu_idx": (cap, np.uint32), "gpu_total": (cap, np.int64),
              "gpu_vals": (cap * 4, np.int64), "counts": (N_COUNTS, np.int64)}
    out = {}
    for k, (m, dt) in shapes.items():
        if pinned:
            t = torch.empty(m * np.dtype(dt).itemsize, dtype=torch.uint8).pin_memory()
            out[k] = t.numpy().view(dt)
            out["_pin_" + k] = t
        else:
            out[k] = np.zeros(m, dt)
    out["defer"] = np.zeros(max(1, defer_cap), _DEFER_DT)
    return out


# ---------------------------------------------------------------------------------------------
# Device/host results -> FleetResult (the bignum escape lives here).
# ---------------------------------------------------------------------------------------------
def _resolve_wide(defer: np.ndarray, rows: np.ndarray) -> Dict[int, Dict[str, int]]:
    """Deferred nodes: add up value texts the device validated but could not hold in int64."""
    wide: Dict[int, Dict[str, int]] = {}
    for rec in defer:
        bd = {}
        for j, k in enumerate(KEYS):
            off, ln = int(rec["off"][j]), int(rec["len"][j])
            if off != 0xFFFFFFFF:
                bd[k] = int(rows[off:off + ln].tobytes().decode("utf-8", "surrogatepass"))
        wide[int(rec["node"])] = bd
    return wide


def finish(flags: np.ndarray, gpu_idx: np.ndarray, gpu_total: np.ndarray, gpu_vals: np.ndarray, counts: np.ndarray,
           defer: np.ndarray, rows: np.ndarray, row_base: int = 0) -> FleetResult:
    """Trim device outputs by the counts block and fold the deferred (bignum) nodes in, in fleet order."""
    if int(counts[C_OUT_DROPPED]) or int(counts[C_DEFER_DROPPED]):
        raise native.GncError(-6, f"output tables too small: {int(counts[C_OUT_DROPPED])} GPU nodes and "
                                  f"{int(counts[C_DEFER_DROPPED])} deferred nodes dropped")
    g, d = int(counts[C_GPU_NODES]), int(counts[C_DEFERRED])
    idx = gpu_idx[:g].astype(np.int64)
    vals = np.asarray(gpu_vals).reshape(-1, 4)[:g]
    totals: List[int] = [int(x) for x in gpu_total[:g]] if g < 4096 else gpu_total[:g].tolist()
    key_tot = [key_total(counts, k) for k in range(4)]
    wide = _resolve_wide(defer[:d], rows) if d else {}
    flags = flags.copy()
    add = sorted(n for n, bd in wide.items() if sum(bd.values()) > 0)          # check-gpu-node.py:222
    if add:
        pos = np.searchsorted(idx, np.asarray(add, dtype=np.int64))
        idx = np.insert(idx, pos, add)
        vals = np.insert(vals, pos, 0, axis=0)
        for p_, n in sorted(zip(pos.tolist(), add), reverse=True):
            totals.insert(p_, sum(wide[n].values()))
        for n in add:
            flags[n - row_base] |= F_GPU
            for j, k in enumerate(KEYS):
                key_tot[j] += wide[n].get(k, 0)
    return FleetResult(len(flags), flags, idx, totals, vals, key_tot, wide, counts.copy(), row_base)
