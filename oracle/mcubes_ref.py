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
        lib.mc_ref_free.argtypes = [ctypes.c_void_p]
        _lib = lib
    return _lib


def marching_cubes(volume, isovalue, return_cases=False):
    lib = _load()
    vol = np.ascontiguousarray(volume, dtype=np.float32)
    nx, ny, nz = vol.shape
    pv, pf = ctypes.POINTER(ctypes.c_double)(), ctypes.POINTER(ctypes.c_int64)()
    nv, nf = ctypes.c_int64(), ctypes.c_int64()
    cases = np.zeros(max((nx - 1) * (ny - 1) * (nz - 1), 1), dtype=np.uint8) if return_cases else None
    lib.mc_ref_run(vol.ctypes.data, nx, ny, nz, float(isovalue), ctypes.byref(pv), ctypes.byref(nv), ctypes.byref(pf),
                   ctypes.byref(nf), cases.ctypes.data if return_cases else None)
    verts = np.ctypeslib.as_array(pv, shape=(nv.value, 3)).copy() if nv.value else np.zeros((0, 3))
    faces = np.ctypeslib.as_array(pf, shape=(nf.value, 3)).copy() if nf.value else np.zeros((0, 3), dtype=np.int64)
    lib.mc_ref_free(pv), lib.mc_ref_free(pf)
    if return_cases:
        return verts, faces, cases.reshape(nx - 1, ny - 1, nz - 1)
    return verts, faces
