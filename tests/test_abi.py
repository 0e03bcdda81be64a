"""CPU: the C-ABI library loads, exports every entry point that include/na_b200.h declares, and the ctypes
binding covers all of them (no compute calls: there is no GPU here)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_entry_points():
    text = open(os.path.join(ROOT, "include", "na_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(na_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_expected_surface():
    names = declared_entry_points()
    for must in ("na_hashgrid_fwd", "na_hashgrid_bwd", "na_hashgrid_taps_fwd", "na_hashgrid_taps_bwd", "na_sdf_mlp_fwd", "na_sdf_head_fwd_planes",
                 "na_sdf_mlp_bwd", "na_mlp_chain_fwd", "na_neus_render_fwd", "na_neus_render_bwd", "na_alpha_weights_fwd",
                 "na_sample_pdf", "na_neus_upsample", "na_merge_sorted", "na_mc_count", "na_mc_emit", "na_adamw_step",
                 "na_weight_norm_fwd", "na_weight_norm_bwd", "na_last_error", "na_abi_version"):
        assert must in names, must


def test_library_exports_every_declared_symbol(na_lib):
    from neuralangelo_b200 import _lib
    raw = ctypes.CDLL(_lib.LIB_PATH)
    missing = [n for n in declared_entry_points() if not hasattr(raw, n)]
    assert not missing, f"declared in include/na_b200.h but not exported: {missing}"


def test_ctypes_binding_covers_every_declared_symbol(na_lib):
    from neuralangelo_b200 import _lib
    bound = set(_lib.exported_signatures())
    unbound = [n for n in declared_entry_points() if n not in bound]
    assert not unbound, f"declared but not bound in neuralangelo_b200/_lib.py: {unbound}"


def test_abi_version_and_error_string(na_lib):
    assert na_lib.na_abi_version() >= 1
    assert isinstance(na_lib.na_last_error(), bytes)
    # host-only entry point: argument validation without a GPU
    from neuralangelo_b200 import _lib
    desc = _lib.HashGridDesc()
    assert na_lib.na_hashgrid_make_desc(64, 8, 22, 32, 1.3, ctypes.byref(desc)) == -1
    assert b"n_levels" in na_lib.na_last_error()
    assert na_lib.na_mc_scratch_bytes(129, 129, 129) > 128 ** 3 * 5
    assert na_lib.na_mc_static_max_cells() == 64 ** 3


def test_product_refuses_cpu_tensors(na_lib):
    """No CPU fallback: the wrappers raise instead of computing on the host."""
    import pytest
    import torch
    from neuralangelo_b200 import _lib, render
    with pytest.raises(_lib.NaError):
        render.alpha_compositing_weights(torch.rand(1, 2, 3))


def test_build_fingerprint_does_not_depend_on_the_mount_path(tmp_path):
    """The library built in the authoring container must be recognised as current when the tree is mounted elsewhere
    (the GPU box): the fingerprint hashes file names and contents, not absolute paths.  (Ranks of one launch used to
    race to rebuild it; build() now also takes an exclusive lock and replaces the .so atomically.)"""
    import importlib.util
    import shutil
    dst = tmp_path / "elsewhere"
    shutil.copytree(os.path.join(ROOT, "neuralangelo_b200"), dst / "neuralangelo_b200",
                    ignore=shutil.ignore_patterns("*.so", "*.o", "build", "__pycache__"))
    shutil.copytree(os.path.join(ROOT, "include"), dst / "include")
    spec = importlib.util.spec_from_file_location("build_elsewhere", dst / "neuralangelo_b200" / "build.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    from neuralangelo_b200 import build
    assert mod._fingerprint() == build._fingerprint()


def test_plane_entry_point_validates_its_envelope(na_lib):
    """na_sdf_head_fwd_planes rejects shapes outside the pipelined kernel's envelope with a message (host-side checks
    run before any device work)."""
    import ctypes as C
    fake = C.c_void_p(256)   # non-null, 16-byte aligned; never dereferenced: validation fails first
    rc = na_lib.na_sdf_head_fwd_planes(fake, 8 * 128, 128, 131, 192, fake, fake, fake, 144, None, None, 100.0, None, None, 0, None)
    assert rc != 0 and b"hidden width" in na_lib.na_last_error()
    rc = na_lib.na_sdf_head_fwd_planes(fake, 8 * 128, 128, 19, 256, fake, fake, fake, 32, None, None, 100.0, None, None, 0, None)
    assert rc != 0 and b"k0=19" in na_lib.na_last_error()
    rc = na_lib.na_sdf_head_fwd_planes(fake, 8 * 100, 128, 131, 256, fake, fake, fake, 144, None, None, 100.0, None, None, 0, None)
    assert rc != 0 and b"chunk_stride" in na_lib.na_last_error()
    assert na_lib.na_sdf_head_fwd_planes(fake, 8 * 128, 0, 131, 256, fake, fake, fake, 144, None, None, 100.0, None, None, 0, None) == 0


def test_new_entry_points_validate_their_envelopes_without_a_gpu(na_lib):
    """Argument checks of the round-2 entry points run before any CUDA call: wrong shapes / null pointers come back as
    NA_ERR_INVALID with a message, never as a crash."""
    from neuralangelo_b200 import _lib, hashgrid
    L = _lib.lib()
    desc = hashgrid.make_desc(16, 8, 22, 32, 1.3195079107728942)
    one = ctypes.c_void_p(256)   # a non-null, 256-byte aligned "pointer" that is never dereferenced: validation fails first
    rc = L.na_lattice_sdf(ctypes.byref(desc), one, one, one, one, 4, 4, 4, _lib.UnitMap(-2.0, 0.25), 16, one, one, 128, 256,
                          one, one, 1.0, one, None)
    assert rc == -1 and b"k0_pad" in L.na_last_error()
    small = hashgrid.make_desc(4, 8, 10, 4, 2.0)
    rc = L.na_lattice_sdf(ctypes.byref(small), one, one, one, one, 4, 4, 4, _lib.UnitMap(-2.0, 0.25), 4, one, one, 144, 256,
                          one, one, 1.0, one, None)
    assert rc == -1 and b"levels" in L.na_last_error()
    rc = L.na_sdf_head_ws(one, 64, one, 131, 8, one, one, 144, 256, one, one, 1.0, one, None)
    assert rc == -1 and b"exactly one" in L.na_last_error()
    rc = L.na_pack_split_f16(one, 278, 250, 250, 278, 0, 0, 1.0, None, 288, one, one, None)
    assert rc == -1 and b"multiple of 8" in L.na_last_error()
    rc = L.na_ray_setup(one, one, one, 1, 8, 0, one, one, one, None, None, None, None)
    assert rc == -1 and b"ray_setup" in L.na_last_error()
    # launch count of the large-batch backward (mirrors launch_bwd): dense levels interleaved + whole levels + 4 x 2 passes
    assert hashgrid.bwd_launch_count(desc, 1 << 22, 7, 16) == 10
    assert hashgrid.bwd_launch_count(desc, 1 << 22, 1, 16) == 2
    assert hashgrid.bwd_launch_count(desc, 1 << 16, 7, 16) == 1
    assert hashgrid.bwd_launch_count(desc, 1 << 22, 7, 4) == 1
