"""ctypes front end of oracle/mcubes_ref.c -- TEST INFRASTRUCTURE ONLY.
`marching_cubes(volume, isovalue)` has the call shape of `mcubes.marching_cubes` (reference call site
projects/neuralangelo/utils/mesh.py:121): float32/float64 volume [X,Y,Z] -> (vertices float64 [nV,3] in
index units, triangles int64 [nF,3])."""
import ctypes

import numpy as np

from . import build_oracle

_lib = None


def _load():
    global _lib
    if _lib is None:
        lib = ctypes.CDLL(build_oracle.build())
        lib.mc_ref_run.restype = ctypes.c_int
        lib.mc_ref_run.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_double,
                                   ctypes.POINTER(ctypes.POINTER(ctypes.c_double)), ctypes.POINTER(ctypes.c_int64),
                                   ctypes.POINTER(ctypes.POINTER(ctypes.c_int64)), ctypes.POINTER(ctypes.c_int64),
                                   ctypes.c_void_p]
        lib.mc_ref_run_mode.restype = ctypes.c_int
        lib.mc_ref_run_mode.argtypes = lib.mc_ref_run.argtypes + [ctypes.c_int]
        lib.mc_ref_free.argtypes = [ctypes.c_void_p]
        _lib = lib
    return _lib


def marching_cubes(volume, isovalue, return_cases=False, dup_low_faces=False):
    """dup_low_faces: the other plausible upstream convention for lattice edges on the low faces of the volume
    (duplicate vertices there; see the header of mcubes_ref.c).  Default = one vertex per crossed edge."""
    lib = _load()
    vol = np.ascontiguousarray(volume, dtype=np.float32)
    nx, ny, nz = vol.shape
    pv, pf = ctypes.POINTER(ctypes.c_double)(), ctypes.POINTER(ctypes.c_int64)()
    nv, nf = ctypes.c_int64(), ctypes.c_int64()
    cases = np.zeros(max((nx - 1) * (ny - 1) * (nz - 1), 1), dtype=np.uint8) if return_cases else None
    lib.mc_ref_run_mode(vol.ctypes.data, nx, ny, nz, float(isovalue), ctypes.byref(pv), ctypes.byref(nv),
                        ctypes.byref(pf), ctypes.byref(nf), cases.ctypes.data if return_cases else None,
                        1 if dup_low_faces else 0)
    verts = np.ctypeslib.as_array(pv, shape=(nv.value, 3)).copy() if nv.value else np.zeros((0, 3))
    faces = np.ctypeslib.as_array(pf, shape=(nf.value, 3)).copy() if nf.value else np.zeros((0, 3), dtype=np.int64)
    lib.mc_ref_free(pv), lib.mc_ref_free(pf)
    if return_cases:
        return verts, faces, cases.reshape(nx - 1, ny - 1, nz - 1)
    return verts, faces


def canonical_faces(verts, faces, decimals=7):
    """Order- and numbering-independent form of a triangle mesh (SURVEY.md section 8c "identical face set after
    canonical relabelling"): every corner is replaced by a key of the lattice edge that owns it -- its position, which
    for a marching-cubes vertex is (two lattice integers, one interpolated coordinate), rounded to `decimals` places
    because the SAME lattice edge interpolated from its other end (a different creating cell) may differ in the last
    ulp -- each triangle is rotated (orientation kept) so that its smallest key comes first, and the triangles are
    sorted.  Two extractions agree as surfaces iff these arrays are equal, however vertices were numbered, ordered
    or duplicated.  Returns int64 [nF, 3] of ranks into the sorted list of distinct keys, plus that list [nU, 3]."""
    verts = np.asarray(verts, dtype=np.float64).reshape(-1, 3)
    faces = np.asarray(faces, dtype=np.int64).reshape(-1, 3)
    if len(faces) == 0:
        return np.zeros((0, 3), dtype=np.int64), np.zeros((0, 3))
    keys = np.round(verts, decimals) + 0.0                              # (+0.0 folds -0.0 into 0.0)
    uniq, inverse = np.unique(keys, axis=0, return_inverse=True)       # lexicographic order of positions
    f = inverse.reshape(-1)[faces]
    first = np.argmin(f, axis=1)
    rows = np.arange(len(f))
    f = np.stack([f[rows, first], f[rows, (first + 1) % 3], f[rows, (first + 2) % 3]], axis=1)
    order = np.lexsort((f[:, 2], f[:, 1], f[:, 0]))
    return f[order], uniq
