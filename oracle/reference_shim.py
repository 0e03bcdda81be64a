"""TEST INFRASTRUCTURE ONLY (see oracle/README.md).

Import shim for the *unmodified* reference under /root/reference so that its own
pure-torch hot-path functions run on CPU in the authoring container.  It exists
to (a) validate the restatements in oracle/*.py and (b) generate the golden
fixtures under tests/golden/ (oracle/make_golden.py).  /root/reference does not
exist on the GPU box, so nothing that runs there may import this module.

What is stubbed: tinycudann, mcubes, trimesh, termcolor are
absent from this image; the reference imports them at module scope
(projects/neuralangelo/utils/modules.py:16, utils/mesh.py:14-15,
imaginaire/utils/termcolor.py:15).  nerf_util.sample_dists has a hard-coded
device="cuda" default (projects/nerf/utils/nerf_util.py:20) that model.py:234
relies on; it is re-bound to "cpu".
"""
import functools
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("NA_REFERENCE_ROOT", "/root/reference")


def available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "projects", "neuralangelo"))


def install(tcnn_module=None):
    """Put the reference on sys.path with stub modules; returns the `projects` namespace."""
    if not available():
        raise RuntimeError(f"reference tree not found at {REFERENCE_ROOT}")
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    if "termcolor" not in sys.modules:
        m = types.ModuleType("termcolor")
        m.colored = lambda s, *a, **k: s
        sys.modules["termcolor"] = m
    for name in ("mcubes", "trimesh"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    if tcnn_module is None:
        from . import tcnn_stub as tcnn_module  # pure-torch hash grid (oracle/tcnn_stub.py)
    sys.modules["tinycudann"] = tcnn_module
    mod = sys.modules.get("projects.neuralangelo.utils.modules")
    if mod is not None:
        mod.tcnn = tcnn_module
    import torch
    from projects.nerf.utils import nerf_util
    if not isinstance(nerf_util.sample_dists, functools.partial):
        dev = "cuda" if torch.cuda.is_available() and os.environ.get("NA_REF_DEVICE", "cpu") == "cuda" else "cpu"
        nerf_util.sample_dists = functools.partial(nerf_util.sample_dists, device=dev)
    import projects
    return projects
