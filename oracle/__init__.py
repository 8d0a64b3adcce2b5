"""CPU oracle for the node-classification path -- TEST INFRASTRUCTURE ONLY.

Nothing under `oracle/` is on the product path.  Only `tests/`, `__graft_entry__.smoke()`
and `bench.py`'s CPU-baseline / `--impl reference` legs may import or execute it, and only
as the checker (or the timed CPU baseline), never as the thing shipped.

Parity status: PINNED.  The reference holds no tests or golden vectors
(/root/reference has 7 files, none of them tests), so the pins are outputs of the reference
itself: `oracle/make_golden.py` imports /root/reference/check-gpu-node.py (with the
`k8s_stub` stand-in for the un-installable `kubernetes` package), runs its own functions and
its own `main()` and commits the results under `tests/golden/`.  `tests/test_oracle_*.py`
hold every restatement here (Python and C) to those vectors.
"""
import os
import sys

# the stand-in lives with the tests (tests/k8s_stub): it is test scaffolding, and code outside oracle/ that needs
# V1Node objects (bench.py's seam leg) can put it on sys.path without importing this package
STUB_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "k8s_stub")


def ensure_kubernetes_importable():
    """Put the stub on sys.path unless a real `kubernetes` package is importable."""
    try:
        import kubernetes  # noqa: F401
    except ImportError:
        sys.path.insert(0, STUB_DIR)
        import kubernetes  # noqa: F401
