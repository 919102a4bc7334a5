"""Compile oracle/cpu_ref.c (the C + OpenMP CPU restatement; test/baseline
infrastructure) into oracle/libcpu_ref.so.  gcc only, no other dependencies."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "cpu_ref.c")
OUT = os.path.join(HERE, "libcpu_ref.so")


def build(verbose: bool = False) -> str:
    if os.path.exists(OUT) and os.path.getmtime(OUT) >= os.path.getmtime(SRC):
        return OUT
    # no -march=native: the .so is built in this container and travels to the GPU box
    cmd = ["gcc", "-O3", "-mavx2", "-mfma", "-fopenmp", "-shared", "-fPIC", "-o", OUT, SRC]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("gcc failed on cpu_ref.c:\n" + res.stderr)
    if verbose:
        print("[oracle] built", OUT, file=sys.stderr)
    return OUT


if __name__ == "__main__":
    print(build(verbose=True))
