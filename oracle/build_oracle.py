"""Compiles the C part of the oracle (gcc): oracle/mcubes_ref.c -> oracle/_build/libmcubes_ref.so.
TEST INFRASTRUCTURE: building the checker is not using it.  Strict IEEE evaluation (-ffp-contract=off)
so that the GPU kernel, which rounds every double operation individually, can be compared bit for bit.

Reference sources: none are compiled.  The reference's native code on this path (tiny-cuda-nn, PyMCubes)
is not vendored under the reference tree, so there is nothing to build into oracle/_ref/.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
OUT_DIR = os.path.join(HERE, "_build")
LIB = os.path.join(OUT_DIR, "libmcubes_ref.so")


def build(force=False):
    src = os.path.join(HERE, "mcubes_ref.c")
    hdr = os.path.join(ROOT, "include", "na_mc_tables.h")
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= max(os.path.getmtime(src), os.path.getmtime(hdr)):
        return LIB
    gcc = shutil.which("gcc")
    if gcc is None:
        raise RuntimeError("gcc not found: cannot build the marching-cubes oracle")
    os.makedirs(OUT_DIR, exist_ok=True)
    cmd = [gcc, "-O2", "-fPIC", "-shared", "-std=c11", "-ffp-contract=off", "-I", os.path.join(ROOT, "include"), src,
           "-o", LIB]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stderr[-4000:])
        raise RuntimeError("gcc failed on oracle/mcubes_ref.c")
    return LIB


if __name__ == "__main__":
    print("built", build(force="--force" in sys.argv))
