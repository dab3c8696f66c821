"""Compile libllie.so for sm_100a in-tree (nvcc cross-compiles without a GPU)."""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libllie.so")


def build(force: bool = False, verbose: bool = False) -> str:
    """Runs `make` in csrc/ (arch flags: -gencode arch=compute_100a,code=sm_100a -lineinfo)."""
    cmd = ["make", "-C", CSRC, "-j", str(min(8, os.cpu_count() or 1))]
    if force:
        subprocess.run(["make", "-C", CSRC, "clean"], check=True, capture_output=not verbose)
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
    if res.returncode != 0:
        raise RuntimeError("building libllie.so failed (see output above)")
    assert os.path.exists(LIB), LIB
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
