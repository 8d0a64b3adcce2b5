"""CPU-side checks of the C ABI: the library loads, exports what include/gnc.h declares, and the
Unicode tables compiled into it are this interpreter's."""
import ctypes
import os
import re

import pytest

from k8s_gpu_node_checker_b200 import native
from oracle import cgn_oracle as o

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(native.LIB_PATH):
        native.build()
    return native.lib()


def test_header_and_library_agree(lib):
    hdr = open(os.path.join(ROOT, "include", "gnc.h")).read()
    declared = set(re.findall(r"\b(gnc_[a-z_]+)\s*\(", hdr))
    assert declared == set(native.EXPORTS)
    for name in declared:
        assert getattr(lib, name) is not None
    assert int(re.search(r"#define GNC_ABI_VERSION (\d+)", hdr).group(1)) == lib.gnc_abi_version() == native.ABI_VERSION


def test_struct_layouts_match_header():
    assert ctypes.sizeof(native.Batch) == 32
    assert ctypes.sizeof(native.DeferRec) == 36
    assert ctypes.sizeof(native.Result) == 56


def test_unicode_tables_are_this_interpreters(lib):
    runs, spaces = native.unicode_tables()
    assert runs == o.unicode_decimal_runs()
    assert spaces == o.unicode_wide_spaces()


def test_no_cpu_fallback_in_product():
    """The product package never imports the oracle."""
    pkg = os.path.join(ROOT, "k8s-gpu-node-checker_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", txt, re.M), f
                assert "pnl_oracle" not in txt, f


def test_traffic_capture_names_the_kernel_variant_the_bench_runs(lib):
    """bench.py reports roofline.traffic from profiles/classify_traffic.json only when that capture is of the
    kernel variant gnc_classify picks for the workload: here the file is held to the library's own choice."""
    import json
    with open(os.path.join(ROOT, "profiles", "classify_traffic.json")) as f:
        t = json.load(f)
    assert lib.gnc_variant_for(t["rows_bytes"], t["nodes"]).decode() == t["variant"]
    assert t["kernel"] == "gnc::classify_kernel" and t["dram_bytes_read"] > t["rows_bytes"] * 0.9
    assert lib.gnc_variant_for(10 ** 9, 10 ** 6).decode() == "t32"       # 1000-byte rows: the smallest tile
    assert lib.gnc_variant_for(300 * 10 ** 6, 10 ** 6).decode() == "t192g2"
