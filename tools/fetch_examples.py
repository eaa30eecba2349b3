#!/usr/bin/env python
"""Copies the reference's examples/*.qasm (151 circuits, 22 MB) into
tests/golden/circuits_full/ - git-ignored input data of BASELINE.json's
config 3 that travels to the GPU box with the snapshot, like oracle/_ref.
Run where /root/reference exists; bench.py --workload c3 and the config-3
tests use the directory when it is there, else the committed 22-circuit subset
(tests/golden/circuits).  Data only: no reference code is copied."""
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(os.environ.get("QMAP_REFERENCE", "/root/reference"), "examples")
DST = os.path.join(ROOT, "tests", "golden", "circuits_full")


def main() -> int:
    if not os.path.isdir(SRC):
        print(f"{SRC} not found: nothing copied", file=sys.stderr)
        return 0
    os.makedirs(DST, exist_ok=True)
    n = 0
    for f in sorted(os.listdir(SRC)):
        if f.endswith(".qasm"):
            shutil.copyfile(os.path.join(SRC, f), os.path.join(DST, f))
            n += 1
    print(f"{n} circuits -> {DST}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
