"""ORACLE — test infrastructure only.  Snapshots the reference's own Python modules for the hot path into oracle/_ref/
(git-ignored; it travels to the GPU box with the gpurun snapshot like a built .so) so that `bench.py --impl reference` and
the `cpu_baseline` legs can time the REFERENCE'S OWN code on the box's host cores (kind "reference") instead of the
oracle port.  Nothing is copied into the tracked tree; run by __graft_entry__.build() whenever /root/reference exists.

    python oracle/snapshot_ref.py [reference_root]
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref")


def snapshot(src_root="/root/reference"):
    src = os.path.join(src_root, "m3ae", "modules")
    if not os.path.isdir(src):
        return False
    dst = os.path.join(DST, "m3ae", "modules")
    if os.path.isdir(DST):
        shutil.rmtree(DST)
    shutil.copytree(src, dst, ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
    for extra in ("label2ans.json",):
        p = os.path.join(src_root, extra)
        if os.path.exists(p):
            shutil.copy(p, os.path.join(DST, extra))
    with open(os.path.join(DST, "README"), "w") as f:
        f.write("Unmodified copy of <reference>/m3ae/modules made by oracle/snapshot_ref.py; git-ignored, never edited.\n")
    return True


if __name__ == "__main__":
    ok = snapshot(sys.argv[1] if len(sys.argv) > 1 else "/root/reference")
    print("oracle/_ref", "written" if ok else "skipped (no reference tree)")
