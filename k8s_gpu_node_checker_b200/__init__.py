"""Import name for the package whose sources live in `k8s-gpu-node-checker_b200/`.

The directory the project layout names has hyphens, which Python cannot import; this shim
extends `__path__` so `k8s_gpu_node_checker_b200.<module>` resolves to the files there.
"""
import os as _os

_SRC = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                     "k8s-gpu-node-checker_b200")
__path__.insert(0, _SRC)
SOURCE_DIR = _SRC
