#!/usr/bin/env python
"""Stages the reference's own CALLERS of the rasterizer plugin, byte for byte, into baseline/_ref/htgs_callers/
(git-ignored like oracle/_ref; it travels to the GPU box with the snapshot) so that tests/test_gpu_reference_callers.py
can import and run them UNCHANGED against this repository's diff_gaussian_rasterization / simple_knn.

What is staged: the Python packages on the plugin's caller side -- gaussian_renderer/, scene/, utils/, arguments/ --
`.py` files only.  Nothing is edited; third-party packages the image lacks (lietorch, plyfile, matplotlib, open3d, ...)
are replaced by import stubs inside the test, not here.  Run where /root/reference is mounted:

    python tools/stage_reference_callers.py            # also run by __graft_entry__.build()
"""
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("HTGS_REFERENCE", "/root/reference")
DST = os.path.join(ROOT, "baseline", "_ref", "htgs_callers")
PACKAGES = ("gaussian_renderer", "scene", "utils", "arguments", "trainer")


def stage():
    if not os.path.isdir(os.path.join(REF, "gaussian_renderer")):
        print(f"reference tree absent at {REF}: keeping whatever is already staged at {DST}")
        return False
    if os.path.isdir(DST):
        shutil.rmtree(DST)
    n = 0
    for pkg in PACKAGES:
        for base, _dirs, files in os.walk(os.path.join(REF, pkg)):
            for f in files:
                if not f.endswith(".py"):
                    continue
                src = os.path.join(base, f)
                dst = os.path.join(DST, os.path.relpath(src, REF))
                os.makedirs(os.path.dirname(dst), exist_ok=True)
                shutil.copyfile(src, dst)
                n += 1
    print(f"staged {n} files into {DST}")
    return True


if __name__ == "__main__":
    sys.exit(0 if stage() or os.path.isdir(DST) else 1)
