"""Load the UNMODIFIED reference script as a module.

`check-gpu-node.py` has hyphens in its name, so it is loaded by path: from /root/reference in the
build container, else from the byte-compiled copy `oracle/build_ref.py` leaves in oracle/_ref/ (the
only form in which the reference reaches the GPU box).  Callers: `oracle/make_golden.py`, the
`needs_reference` tests, and `bench.py`'s CPU legs (the timed CPU baseline) -- never the product.
"""
import importlib.machinery
import importlib.util
import os

from . import ensure_kubernetes_importable
from .build_ref import REF_PYC, REFERENCE_SCRIPT


def reference_path():
    """The script in /root/reference if present, else the compiled copy, else None."""
    if os.path.isfile(REFERENCE_SCRIPT):
        return REFERENCE_SCRIPT
    if os.path.isfile(REF_PYC):
        return REF_PYC
    return None


def reference_available() -> bool:
    return reference_path() is not None


def load_reference(path: str | None = None):
    ensure_kubernetes_importable()
    path = path or reference_path()
    if path is None:
        raise FileNotFoundError("neither /root/reference/check-gpu-node.py nor oracle/_ref/check_gpu_node.bin exists")
    loader = (importlib.machinery.SourcelessFileLoader if path.endswith((".pyc", ".bin")) else importlib.machinery.SourceFileLoader)(
        "cgn_reference", path)
    spec = importlib.util.spec_from_loader("cgn_reference", loader)
    mod = importlib.util.module_from_spec(spec)
    loader.exec_module(mod)
    return mod
