"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol
include/radnerf_b200.h declares; argument validation works without a GPU."""
import ctypes

import pytest

from radnerf_b200 import _lib


def test_library_exports_every_declared_symbol():
    names = _lib.declared_symbols()
    assert len(names) >= 15
    lib = _lib.lib()
    for n in names:
        assert hasattr(lib, n), n


def test_abi_version():
    assert _lib.lib().rnb_abi_version() == 1


def test_bad_argument_is_reported_not_crashed():
    lib = _lib.lib()
    rc = lib.rnb_morton3D(ctypes.c_void_p(0), ctypes.c_int64(-1), ctypes.c_void_p(0), ctypes.c_void_p(0))
    assert rc == -1
    assert b"rnb_morton3D" in lib.rnb_last_error()
    # empty input is a no-op, as for the reference (zero-sized launch)
    assert lib.rnb_morton3D(ctypes.c_void_p(0), ctypes.c_int64(0), ctypes.c_void_p(0), ctypes.c_void_p(0)) == 0


def test_vren_shim_has_the_reference_names():
    from radnerf_b200 import vren
    for n in ["ray_aabb_intersect", "ray_sphere_intersect", "morton3D", "morton3D_invert", "packbits",
              "raymarching_train", "raymarching_test", "composite_train_fw", "composite_train_bw",
              "composite_test_fw", "distortion_loss_fw", "distortion_loss_bw"]:
        assert callable(getattr(vren, n))


def test_check_input_matches_reference_contract():
    import torch
    from radnerf_b200 import vren
    with pytest.raises(RuntimeError, match="must be a CUDA tensor"):
        vren.morton3D(torch.zeros(4, 3, dtype=torch.int32))
