"""Run the reference CLI with the B200 classification path behind it.

    python -m k8s_gpu_node_checker_b200.cli --reference /path/to/check-gpu-node.py [its own flags ...]

Nothing of the reference's command line is re-implemented: its script is loaded as a module, its
`list_gpu_nodes` (check-gpu-node.py:215-226) is replaced by `api.list_gpu_nodes`, and its own `main()`
(:314-327) runs -- so all seven flags, the kubeconfig lookup, the Slack gating and message, the table,
the JSON document and the exit codes (:137-157, :229-293, :298-311) are the reference's, byte for byte.
Two additive options are consumed here and never reach the reference's parser:

    --reference PATH        the reference script (default: $GNC_REFERENCE_SCRIPT)
    --engine b200|reference which list_gpu_nodes to use (default b200; `reference` leaves it untouched)
"""
from __future__ import annotations

import importlib.util
import os
import sys
from typing import List, Optional


def load_reference(path: str):
    spec = importlib.util.spec_from_file_location("check_gpu_node_reference", path)
    if spec is None or spec.loader is None:
        raise FileNotFoundError(path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _take(argv: List[str], flag: str, default: Optional[str]) -> Optional[str]:
    """Remove `flag VALUE` / `flag=VALUE` from argv and return VALUE."""
    for i, a in enumerate(argv):
        if a == flag and i + 1 < len(argv):
            v = argv[i + 1]
            del argv[i:i + 2]
            return v
        if a.startswith(flag + "="):
            del argv[i]
            return a.split("=", 1)[1]
    return default


def main(argv: Optional[List[str]] = None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    ref_path = _take(argv, "--reference", os.environ.get("GNC_REFERENCE_SCRIPT"))
    engine = _take(argv, "--engine", "b200")
    if not ref_path:
        print("error: --reference PATH (or GNC_REFERENCE_SCRIPT) must name check-gpu-node.py", file=sys.stderr)
        return 1
    if engine not in ("b200", "reference"):
        print(f"error: --engine must be b200 or reference, not {engine!r}", file=sys.stderr)
        return 1
    ref = load_reference(ref_path)
    if engine == "b200":
        from . import api
        ref.list_gpu_nodes = api.list_gpu_nodes          # same signature, same return value
    old = sys.argv
    sys.argv = [ref_path] + argv
    try:
        return ref.main()
    finally:
        sys.argv = old


if __name__ == "__main__":
    try:
        from dotenv import load_dotenv                    # the reference does this before main() (:331)
        load_dotenv()
    except ImportError:
        pass
    sys.exit(main())
