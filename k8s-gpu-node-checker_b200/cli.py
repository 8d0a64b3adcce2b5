"""Run the reference CLI with the B200 classification path behind it.

    python -m k8s_gpu_node_checker_b200.cli --reference /path/to/check-gpu-node.py [its own flags ...]

Nothing of the reference's command line is re-implemented: its script is loaded as a module, its
`list_gpu_nodes` (check-gpu-node.py:215-226) is replaced by `api.list_gpu_nodes`, and its own `main()`
(:314-327) runs -- so all seven flags, the kubeconfig lookup, the Slack gating and message, the table,
the JSON document and the exit codes (:137-157, :229-293, :298-311) are the reference's, byte for byte.
With the b200 engine the `--json` document (:279) is written by the native emitter straight from the
device's decisions (`api.summary_json_from_result`, byte-identical to the reference's `json.dumps` text):
the reference's own `json.dumps(payload, ensure_ascii=False, indent=2)` call is answered from it when the
payload is the one `list_gpu_nodes` just produced, and passed through untouched otherwise.
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
    """The reference script as a module: its source, or a byte-compiled copy of it (.pyc content under any name)."""
    import importlib.machinery
    if not os.path.isfile(path):
        raise FileNotFoundError(path)
    with open(path, "rb") as f:
        compiled = f.read(4) == importlib.util.MAGIC_NUMBER
    loader = (importlib.machinery.SourcelessFileLoader if compiled else importlib.machinery.SourceFileLoader)(
        "check_gpu_node_reference", path)
    spec = importlib.util.spec_from_loader("check_gpu_node_reference", loader)
    mod = importlib.util.module_from_spec(spec)
    loader.exec_module(mod)
    return mod


class _JsonForSummary:
    """Stands in for the `json` module inside the loaded reference: every attribute is the real module's, and
    `dumps` is too, except for the one call of check-gpu-node.py:279 -- same payload object, same arguments --
    whose text the native writer produces from the device's result without touching the dicts again."""

    def __init__(self, real, api):
        self._real, self._api = real, api

    def __getattr__(self, name):
        return getattr(self._real, name)

    def dumps(self, obj, *args, **kw):
        last = self._api.LastCall
        if (not args and kw == {"ensure_ascii": False, "indent": 2} and isinstance(obj, dict)
                and list(obj) == ["total_nodes", "ready_nodes", "nodes"] and last.gpu_nodes is not None
                and obj["nodes"] is last.gpu_nodes and obj["total_nodes"] == len(last.gpu_nodes)):
            if obj["ready_nodes"] == last.result.ready_nodes:
                return self._api.summary_json_from_result(last.nodes, last.result)
        return self._real.dumps(obj, *args, **kw)


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
        ref.json = _JsonForSummary(ref.json, api)        # the --json text comes from the native writer (:279)
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
