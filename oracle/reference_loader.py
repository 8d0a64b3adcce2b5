"""Load the UNMODIFIED reference script as a module (this container only).

`check-gpu-node.py` has hyphens in its name, so it is loaded by path.  /root/reference does
not exist on the GPU box: only `oracle/make_golden.py` and the `needs_reference` tests call
this, never `-m gpu` tests, `smoke()` or `bench.py`.
"""
import importlib.util
import os

from . import ensure_kubernetes_importable

REFERENCE_SCRIPT = "/root/reference/check-gpu-node.py"


def reference_available() -> bool:
    return os.path.isfile(REFERENCE_SCRIPT)


def load_reference():
    ensure_kubernetes_importable()
    spec = importlib.util.spec_from_file_location("cgn_reference", REFERENCE_SCRIPT)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod
