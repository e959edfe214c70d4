"""Build libshipsim.so in-tree with nvcc for sm_100a (the built .so travels to the GPU box with the
repo snapshot; it is git-ignored).  `python -m sac_maritime_ast_b200.csrc.build [--force]`."""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(HERE)
ROOT = os.path.dirname(PKG)
SRC = os.path.join(HERE, "shipsim.cu")
DEPS = [SRC, os.path.join(HERE, "ship_core.cuh"), os.path.join(HERE, "ship_x2.cuh"), os.path.join(HERE, "env_x2.cuh"), os.path.join(HERE, "env_core.cuh"), os.path.join(HERE, "sac_learner.cuh"),
        os.path.join(ROOT, "include", "shipsim.h")]
OUT = os.path.join(PKG, "libshipsim.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared", "--expt-relaxed-constexpr"]


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and os.path.exists(OUT) and all(os.path.getmtime(OUT) >= os.path.getmtime(d) for d in DEPS):
        return OUT
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", OUT, SRC]
    subprocess.run(cmd, check=True)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
