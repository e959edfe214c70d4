"""Build recipe for the C oracle (TEST INFRASTRUCTURE).  gcc only; no fast-math and no FMA
contraction so the float64 arithmetic is the same expression-by-expression IEEE arithmetic the
reference's Python floats perform.

The reference is pure Python (no C/C++ sources), so there is nothing to compile into
``oracle/_ref/``; the real reference is exercised through ``oracle/ref_harness.py`` in the build
container and frozen into ``tests/golden/`` instead.
"""
from __future__ import annotations

import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(_HERE, "ship_oracle.c")
OUT_DIR = os.path.join(_HERE, "_build")
OUT = os.path.join(OUT_DIR, "libshiporacle.so")


def build_oracle(force: bool = False) -> str:
    os.makedirs(OUT_DIR, exist_ok=True)
    deps = [SRC, os.path.join(_HERE, "ship_oracle.h")]
    if not force and os.path.exists(OUT) and all(os.path.getmtime(OUT) >= os.path.getmtime(d) for d in deps):
        return OUT
    cmd = ["gcc", "-O2", "-fPIC", "-shared", "-fopenmp", "-ffp-contract=off", "-Wall", "-o", OUT, SRC, "-lm"]
    subprocess.run(cmd, check=True)
    return OUT


if __name__ == "__main__":
    print(build_oracle(force=True))
