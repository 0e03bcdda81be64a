"""Build recipe for libna_b200.so (explicit nvcc, sm_100a only, in-tree output).

The library is plain CUDA C++ behind a C ABI (include/na_b200.h): no torch headers, no
pybind.  It is built in-tree so that the .so travels with the repository snapshot to
the GPU box; a JIT cache would not.
"""
import hashlib
import os
import shutil
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
REPO_ROOT = os.path.dirname(PKG_DIR)
CSRC = os.path.join(PKG_DIR, "csrc")
INCLUDE = os.path.join(REPO_ROOT, "include")
LIB_PATH = os.path.join(PKG_DIR, "libna_b200.so")
STAMP_PATH = os.path.join(PKG_DIR, ".libna_b200.stamp")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC,-O3,-Wall",
    "-Xptxas", "-v",
] + os.environ.get("NA_NVCC_EXTRA", "").split()   # developer builds only (e.g. -DNA_MLP_PROF)


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def _fingerprint():
    h = hashlib.sha256()
    files = sources() + sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cuh"))
    files += [os.path.join(INCLUDE, f) for f in sorted(os.listdir(INCLUDE))]
    for f in files:
        h.update(os.path.basename(f).encode())   # names, not absolute paths: the tree may be mounted elsewhere
        with open(f, "rb") as fh:
            h.update(fh.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def is_current():
    if not (os.path.exists(LIB_PATH) and os.path.exists(STAMP_PATH)):
        return False
    with open(STAMP_PATH) as fh:
        return fh.read().strip() == _fingerprint()


def build(force=False, verbose=True):
    """Compile every .cu under csrc/ into libna_b200.so.  Returns the library path.  Safe to call from several
    processes at once (the ranks of one launch): an exclusive file lock serialises them and the late comers find
    the library current."""
    if not force and is_current():
        return LIB_PATH
    import fcntl
    obj_dir = os.path.join(PKG_DIR, "build")
    os.makedirs(obj_dir, exist_ok=True)
    with open(os.path.join(obj_dir, ".lock"), "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and is_current():
                return LIB_PATH
            return _build_locked(obj_dir, verbose)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


def _build_locked(obj_dir, verbose):
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found: cannot build libna_b200.so")
    objs = []
    log = []
    for src in sources():
        obj = os.path.join(obj_dir, os.path.basename(src)[:-3] + ".o")
        cmd = [nvcc, *NVCC_FLAGS, "-I", INCLUDE, "-I", CSRC, "-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        log.append(r.stderr)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError(f"nvcc failed on {src}")
        objs.append(obj)
    tmp_lib = LIB_PATH + ".tmp"
    cmd = [nvcc, "-shared", "-o", tmp_lib, *objs, "-lcudart"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("link failed")
    os.replace(tmp_lib, LIB_PATH)
    with open(os.path.join(obj_dir, "ptxas.log"), "w") as fh:
        fh.write("\n".join(log))
    with open(STAMP_PATH, "w") as fh:
        fh.write(_fingerprint())
    if verbose:
        print(f"built {LIB_PATH}")
    return LIB_PATH


if __name__ == "__main__":
    build(force="--force" in sys.argv)
