"""Recipe for oracle/_ref/: the reference's own code, compiled where it lies (TEST INFRASTRUCTURE).

    python -m oracle.build_ref

The reference is one Python script, so "compiling from its sources" is byte-compiling it:
/root/reference/check-gpu-node.py -> oracle/_ref/check_gpu_node.bin (py_compile, unoptimised, the
interpreter of this image; a .pyc by content -- the name avoids the suffix that repository snapshots
leave out).  No source text is copied into the repository; oracle/_ref/ is
git-ignored and travels to the GPU box with the tree like the other built files, which lets
`bench.py`'s CPU legs time the reference's own `list_gpu_nodes` (check-gpu-node.py:215-226) on that
box's host cores and lets the CLI tests run the reference's own `main()` there.  The build container
has /root/reference; the GPU box only ever sees the compiled file.
"""
from __future__ import annotations

import os
import py_compile

HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE_SCRIPT = "/root/reference/check-gpu-node.py"
REF_DIR = os.path.join(HERE, "_ref")
REF_PYC = os.path.join(REF_DIR, "check_gpu_node.bin")


def build(force: bool = False) -> str | None:
    """Byte-compile the reference script into oracle/_ref/.  Returns the path, or None when
    /root/reference is not on this machine (then whatever was built earlier stays in place)."""
    if not os.path.isfile(REFERENCE_SCRIPT):
        return REF_PYC if os.path.isfile(REF_PYC) else None
    os.makedirs(REF_DIR, exist_ok=True)
    if force or not os.path.isfile(REF_PYC) or os.path.getmtime(REF_PYC) < os.path.getmtime(REFERENCE_SCRIPT):
        py_compile.compile(REFERENCE_SCRIPT, cfile=REF_PYC, doraise=True,
                           invalidation_mode=py_compile.PycInvalidationMode.UNCHECKED_HASH)
    return REF_PYC


if __name__ == "__main__":
    print(build(force=True))
