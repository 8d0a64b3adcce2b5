"""PNL-v2 -- the packed node list the device reads (spec: DESIGN.md "Data layout").

One byte stream plus one offset column; every string the reference looks at is kept whole, so
key matching, quantity parsing and the Ready test all happen on the device, not in the packer.

    rows    u8[T]      node rows back to back; node i is rows[row_ptr[i] : row_ptr[i+1]]
    row_ptr u32[N+1]   row_ptr[0] == 0 (or the shard's first byte), non-decreasing, multiples of 4

    row  := n_cap:u8 n_cond:u8  (klen:u8 vlen:u8){n_cap + n_cond}  text  pad
            the length table comes first so a reader steps over entries with integer adds, not
            with one dependent load per entry; `text` is key0 value0 key1 value1 ... in table
            order: first the node's status.capacity (dict order), then its status.conditions as
            (type, status); `pad` is 0..3 zero bytes so that every row starts 4-byte aligned
    xrow := 0xFF 0xFF 0 0  n_cap:u32le n_cond:u32le  (klen:u32le vlen:u32le){n_cap + n_cond}  text  pad
            the extended form, used when a count or a length is >= 255
    Strings are UTF-8 ('surrogatepass').

Marshalling done here (type adaptation only, no classification):
  * `status` falsy or `status.capacity` falsy -> no capacity entries (check-gpu-node.py:183-185)
  * a capacity value that is not a `str` becomes `""` when falsy (the reference skips it, :188)
    and `str(value)` otherwise (the reference applies `str` itself, :192)
  * list elements that are not `V1NodeCondition` instances are left out: the reference's
    `isinstance` test (:176) can never accept them; a non-`str` type/status is stored as `""`
Device buffers must be 16-byte aligned with `rows` padded by >= 32 bytes and `row_ptr` by
>= 4 entries past their last valid element (bulk copies move whole 16-byte units).
"""
from __future__ import annotations

import struct
from dataclasses import dataclass
from typing import Any, Iterable, List, Optional, Sequence, Tuple

import numpy as np

GPU_RESOURCE_KEYS = ("nvidia.com/gpu", "amd.com/gpu", "gpu.intel.com/i915", "intel.com/gpu")

EMPTY_ROW = b"\0\0\0\0"
ROWS_PAD = 32        # bytes readable past T
MAX_ROWS_BYTES = 2 ** 32 - 2 ** 20
ROWPTR_PAD = 4       # entries readable past N+1


def _utf8(s: str) -> bytes:
    return s.encode("utf-8", "surrogatepass")


def marshal_quantity(value: Any) -> bytes:
    """Bytes the device parses for one capacity value (see module docstring)."""
    if isinstance(value, str):
        return _utf8(value)
    try:
        return b"" if not value else _utf8(str(value))
    except Exception:
        return b"?"          # `str()` raised inside the reference's try block -> key dropped


def encode_row(capacity: Optional[Iterable[Tuple[Any, Any]]],
               conditions: Optional[Iterable[Tuple[Any, Any]]]) -> bytes:
    """One row from (key, value) pairs and (type, status) pairs."""
    caps = [(_utf8(k), marshal_quantity(v)) for k, v in capacity or () if isinstance(k, str)]
    # a non-str key cannot equal any of the four str keys: left out
    conds = [(_utf8(t) if isinstance(t, str) else b"", _utf8(s) if isinstance(s, str) else b"")
             for t, s in conditions or ()]
    return encode_raw_row(caps, conds)


def encode_raw_row(caps: Sequence[Tuple[bytes, bytes]], conds: Sequence[Tuple[bytes, bytes]],
                   extended: Optional[bool] = None) -> bytes:
    entries = list(caps) + list(conds)
    if extended is None:
        extended = len(caps) >= 255 or len(conds) >= 255 or any(len(k) >= 255 or len(v) >= 255 for k, v in entries)
    if extended:
        out = bytearray(b"\xff\xff\0\0") + struct.pack("<II", len(caps), len(conds))
        for k, v in entries:
            out += struct.pack("<II", len(k), len(v))
    else:
        out = bytearray((len(caps), len(conds)))
        for k, v in entries:
            out += bytes((len(k), len(v)))
    for k, v in entries:
        out += k
        out += v
    out += b"\0" * (-len(out) % 4)
    return bytes(out)


def _condition_class():
    try:
        from kubernetes.client import V1NodeCondition
    except ImportError as e:          # no silent stand-in: V1Node objects imply the client is there
        raise ImportError("packing V1Node objects needs `kubernetes.client.V1NodeCondition`") from e
    return V1NodeCondition


def node_row(node: Any, cond_cls=None) -> bytes:
    """Row for one V1Node-like object (attribute names of check-gpu-node.py:173-187)."""
    cond_cls = cond_cls or _condition_class()
    st = node.status
    if not st:
        return EMPTY_ROW
    cap = st.capacity.items() if st.capacity else None
    conds = None
    if st.conditions:
        conds = [(c.type, c.status) for c in st.conditions if isinstance(c, cond_cls)]
    return encode_row(cap, conds)


@dataclass
class PackedNodeList:
    rows: np.ndarray        # uint8, length T + ROWS_PAD (zero padded)
    row_ptr: np.ndarray     # uint32, length N + 1 + ROWPTR_PAD (padding repeats T)
    n_nodes: int

    @property
    def n_bytes(self) -> int:
        return int(self.row_ptr[self.n_nodes])

    def row(self, i: int) -> bytes:
        return self.rows[int(self.row_ptr[i]):int(self.row_ptr[i + 1])].tobytes()

    def slice(self, lo: int, hi: int) -> "PackedNodeList":
        """Nodes [lo, hi) as a self-contained list (offsets rebased) -- one rank's shard."""
        b0, b1 = int(self.row_ptr[lo]), int(self.row_ptr[hi])
        return from_parts(self.rows[b0:b1], self.row_ptr[lo:hi + 1].astype(np.int64) - b0)


def from_parts(rows: np.ndarray, row_ptr: np.ndarray) -> PackedNodeList:
    n = len(row_ptr) - 1
    t = int(row_ptr[n]) if n >= 0 else 0
    if t > MAX_ROWS_BYTES:
        raise ValueError("PNL-v2 rows are limited to 4 GiB - 1 MiB per list; split the fleet")
    r = np.zeros(t + ROWS_PAD, dtype=np.uint8)
    r[:t] = rows[:t]
    p = np.full(n + 1 + ROWPTR_PAD, t, dtype=np.uint32)
    p[:n + 1] = row_ptr
    return PackedNodeList(r, p, n)


def from_rows(rows: Sequence[bytes]) -> PackedNodeList:
    lens = np.fromiter((len(r) for r in rows), dtype=np.int64, count=len(rows))
    ptr = np.zeros(len(rows) + 1, dtype=np.int64)
    np.cumsum(lens, out=ptr[1:])
    return from_parts(np.frombuffer(b"".join(rows), dtype=np.uint8), ptr)


def _fastpack():
    """The C packer (csrc/fastpack.c) when it has been built, else None."""
    try:
        from . import _fastpack as fp
        return fp
    except ImportError:
        return None


def pack_nodes(nodes: Optional[Iterable[Any]], use_c: Optional[bool] = None) -> PackedNodeList:
    """V1Node objects (what `api.list_node().items` holds) -> PNL-v2.

    Uses the C extension when it is built (same marshalling, ~20x faster); `use_c=False` forces the
    Python packer, `use_c=True` requires the extension."""
    cls = _condition_class()
    fp = _fastpack() if use_c is not False else None
    if use_c and fp is None:
        raise ImportError("_fastpack is not built; run __graft_entry__.build()")
    if fp is None:
        return from_rows([node_row(n, cls) for n in (nodes or ())])
    rows, ptr, n = fp.pack_nodes(list(nodes or ()), cls)
    return PackedNodeList(np.frombuffer(rows, dtype=np.uint8).copy(), np.frombuffer(ptr, dtype=np.uint32).copy(), n)


# ---------------------------------------------------------------------------------------------
# Reading rows back (for the host side of the bignum escape and for tests).
# ---------------------------------------------------------------------------------------------
def decode_row(row: bytes) -> Tuple[List[Tuple[bytes, bytes]], List[Tuple[bytes, bytes]]]:
    """(capacity entries, condition entries) of one row, as raw byte pairs; ([], []) if malformed."""
    if len(row) < 2:
        return [], []
    if row[0] == 0xFF:
        if len(row) < 12:
            return [], []
        n_cap, n_cond = struct.unpack_from("<II", row, 4)
        n = n_cap + n_cond
        if 12 + 8 * n > len(row):
            return [], []
        lens = struct.unpack_from(f"<{2 * n}I", row, 12)
        pos = 12 + 8 * n
    else:
        n_cap, n_cond = row[0], row[1]
        n = n_cap + n_cond
        if 2 + 2 * n > len(row):
            return [], []
        lens = row[2:2 + 2 * n]
        pos = 2 + 2 * n
    out = []
    for i in range(n):
        kl, vl = lens[2 * i], lens[2 * i + 1]
        if pos + kl + vl > len(row):
            break
        out.append((row[pos:pos + kl], row[pos + kl:pos + kl + vl]))
        pos += kl + vl
    return out[:n_cap], out[n_cap:]
