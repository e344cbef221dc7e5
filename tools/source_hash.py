#!/usr/bin/env python
"""Hash of the sources that determine the CUDA library (csrc/*.cu, csrc/*.cuh, include/lsb200.h, the Makefile).
profiles/ncu_traffic.json is stamped with it at capture time; bench.py refuses a capture whose stamp is not the hash
of the tree it runs from (a traffic figure of another build is not a measurement of this one)."""
import glob
import hashlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def source_hash(root=ROOT):
    files = sorted(glob.glob(os.path.join(root, "lidarshooter_b200", "csrc", "*.cu")) +
                   glob.glob(os.path.join(root, "lidarshooter_b200", "csrc", "*.cuh")) +
                   [os.path.join(root, "include", "lsb200.h"), os.path.join(root, "lidarshooter_b200", "Makefile")])
    h = hashlib.sha256()
    for f in files:
        h.update(os.path.relpath(f, root).encode())
        h.update(open(f, "rb").read())
    return h.hexdigest()[:16]


if __name__ == "__main__":
    sys.stdout.write(source_hash() + "\n")
