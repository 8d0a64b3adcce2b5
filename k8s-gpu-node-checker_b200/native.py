"""ctypes binding of libgnc.so (include/gnc.h).  No CPU stand-in: a missing library is an error."""
from __future__ import annotations

import ctypes
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_PATH = os.path.join(CSRC, "libgnc.so")
ABI_VERSION = 2
N_COUNTS = 16
STATIC_MAX_TILES = 65536

F_GPU, F_READY, F_DEFERRED, F_KEY0, F_MALFORMED = 0x01, 0x02, 0x04, 0x08, 0x80
C_GPU_NODES, C_READY_NODES, C_DEFERRED, C_SCANNED, C_KEY_LO, C_KEY_HI = 0, 1, 2, 3, 4, 8
C_MALFORMED, C_DEFER_DROPPED, C_OUT_DROPPED = 12, 13, 14


class Batch(ctypes.Structure):
    _fields_ = [("rows", ctypes.c_void_p), ("row_ptr", ctypes.c_void_p), ("rows_bytes", ctypes.c_uint64),
                ("n_nodes", ctypes.c_uint32), ("node_base", ctypes.c_uint32)]


class DeferRec(ctypes.Structure):
    _fields_ = [("node", ctypes.c_uint32), ("off", ctypes.c_uint32 * 4), ("len", ctypes.c_uint32 * 4)]


class Result(ctypes.Structure):
    _fields_ = [("flags", ctypes.c_void_p), ("gpu_idx", ctypes.c_void_p), ("gpu_total", ctypes.c_void_p),
                ("gpu_vals", ctypes.c_void_p), ("defer", ctypes.c_void_p), ("counts", ctypes.c_void_p),
                ("out_cap", ctypes.c_uint32), ("defer_cap", ctypes.c_uint32)]


class JsonOut(ctypes.Structure):
    _fields_ = [("rows", ctypes.c_void_p), ("row_ptr", ctypes.c_void_p), ("info", ctypes.c_void_p),
                ("rows_cap", ctypes.c_uint64), ("nodes_cap", ctypes.c_uint32)]


EXPORTS = ("gnc_abi_version", "gnc_create", "gnc_comm_id", "gnc_reserve", "gnc_set_max_str_digits", "gnc_classify",
           "gnc_gather", "gnc_fleet_enable", "gnc_fleet_target", "gnc_fleet_timing", "gnc_classify_host", "gnc_classify_static", "gnc_json_to_pnl", "gnc_launch_count", "gnc_variant_for", "gnc_unicode_tables", "gnc_last_error", "gnc_destroy")


class GncError(RuntimeError):
    def __init__(self, code: int, text: str):
        super().__init__(f"libgnc error {code}: {text}")
        self.code = code


def build(verbose: bool = False) -> str:
    """Compile csrc/ into csrc/libgnc.so for sm_100a (nvcc cross-compiles without a GPU)."""
    out = subprocess.run(["bash", os.path.join(CSRC, "build.sh")], capture_output=True, text=True)
    if verbose or out.returncode:
        print(out.stdout[-4000:], out.stderr[-4000:])
    if out.returncode:
        raise RuntimeError("building libgnc.so failed")
    return LIB_PATH


def build_fastpack() -> str:
    """Compile csrc/fastpack.c (the C packer for V1Node objects) next to the package's Python files."""
    import sysconfig
    out = os.path.join(HERE, "_fastpack" + sysconfig.get_config_var("EXT_SUFFIX"))
    cmd = ["gcc", "-O2", "-fPIC", "-shared", "-Wall", "-I" + sysconfig.get_paths()["include"],
           os.path.join(CSRC, "fastpack.c"), "-o", out]
    subprocess.check_call(cmd)
    return out


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FileNotFoundError(f"{LIB_PATH} is not built; run `python -c 'import __graft_entry__ as g; g.build()'`. "
                                "There is no CPU fallback for the classification path.")
    L = ctypes.CDLL(LIB_PATH)
    vp, i32, u32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_uint32
    L.gnc_abi_version.restype = i32
    L.gnc_create.argtypes = [ctypes.POINTER(vp), i32, i32, i32, vp, ctypes.c_size_t]
    L.gnc_comm_id.argtypes = [vp, ctypes.c_size_t, ctypes.POINTER(ctypes.c_size_t)]
    L.gnc_reserve.argtypes = [vp, u32]
    L.gnc_set_max_str_digits.argtypes = [vp, i32]
    L.gnc_classify.argtypes = [vp, ctypes.POINTER(Batch), ctypes.POINTER(Result), vp]
    L.gnc_gather.argtypes = [vp, ctypes.POINTER(Result), ctypes.POINTER(Result), vp]
    L.gnc_classify_host.argtypes = [vp, ctypes.POINTER(Batch), ctypes.POINTER(Result)]
    L.gnc_fleet_enable.argtypes = [vp, u32, vp]
    L.gnc_fleet_enable.restype = i32
    L.gnc_fleet_target.argtypes = [vp, ctypes.POINTER(Result)]
    L.gnc_fleet_target.restype = i32
    L.gnc_fleet_timing.argtypes = [vp, ctypes.POINTER(ctypes.c_uint64)]
    L.gnc_fleet_timing.restype = i32
    L.gnc_classify_static.argtypes = [i32, ctypes.POINTER(Batch), ctypes.POINTER(Result), i32, vp, ctypes.c_char_p,
                                      ctypes.c_size_t]
    L.gnc_classify_static.restype = i32
    L.gnc_json_to_pnl.argtypes = [vp, vp, ctypes.c_uint64, ctypes.POINTER(JsonOut), vp]
    L.gnc_json_to_pnl.restype = i32
    L.gnc_launch_count.argtypes = [vp]
    L.gnc_launch_count.restype = ctypes.c_uint64
    L.gnc_variant_for.argtypes = [ctypes.c_uint64, u32]
    L.gnc_variant_for.restype = ctypes.c_char_p
    L.gnc_unicode_tables.argtypes = [ctypes.POINTER(ctypes.POINTER(u32)), ctypes.POINTER(i32),
                                     ctypes.POINTER(ctypes.POINTER(u32)), ctypes.POINTER(i32)]
    L.gnc_last_error.argtypes = [vp]
    L.gnc_last_error.restype = ctypes.c_char_p
    L.gnc_destroy.argtypes = [vp]
    L.gnc_destroy.restype = None
    for name in ("gnc_create", "gnc_comm_id", "gnc_reserve", "gnc_set_max_str_digits", "gnc_classify", "gnc_gather",
                 "gnc_classify_host", "gnc_unicode_tables"):
        getattr(L, name).restype = i32
    if L.gnc_abi_version() != ABI_VERSION:
        raise RuntimeError(f"libgnc.so has ABI {L.gnc_abi_version()}, binding expects {ABI_VERSION}; rebuild")
    _lib = L
    return L


def unicode_tables():
    """(decimal-digit run starts, wide spaces) compiled into the library."""
    L = lib()
    pr, ps = ctypes.POINTER(ctypes.c_uint32)(), ctypes.POINTER(ctypes.c_uint32)()
    nr, ns = ctypes.c_int(), ctypes.c_int()
    L.gnc_unicode_tables(ctypes.byref(pr), ctypes.byref(nr), ctypes.byref(ps), ctypes.byref(ns))
    return [pr[i] for i in range(nr.value)], [ps[i] for i in range(ns.value)]
