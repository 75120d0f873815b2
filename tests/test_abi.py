"""CPU-side checks of the boundary: the C-ABI library loads, exports exactly what include/vxorb.h
declares, and fails loudly (no CPU fallback) when there is no GPU."""
import ctypes
import os
import re

import pytest

import visionx_slam_b200 as vx

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "vxorb.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(vxorb_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree():
    assert _declared_symbols() == sorted(vx.EXPORTED_SYMBOLS)


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(vx.LIB_PATH)
    for name in _declared_symbols():
        assert hasattr(lib, name), f"{name} declared in vxorb.h but not exported by libvxorb.so"
    vx.load_library()  # binds argtypes for every symbol


def test_version_and_default_params():
    lib = vx.load_library()
    assert b"sm_100a" in lib.vxorb_version()
    p = vx.Params()
    lib.vxorb_default_params(ctypes.byref(p))
    # reference defaults: orb_extractor.h:11-13, orb_matcher.h:13-16, cv::ORB defaults
    assert (p.n_features, p.n_levels, p.edge_threshold, p.patch_size, p.fast_threshold) == (1000, 8, 31, 31, 20)
    assert abs(p.scale_factor - 1.2) < 1e-6 and abs(p.nn_ratio - 0.8) < 1e-6 and p.min_matches == 50


def test_struct_layouts_match_header():
    assert vx.KEYPOINT_DTYPE.itemsize == 20
    assert vx.MATCH_DTYPE.itemsize == 12
    assert vx.KNN2_DTYPE.itemsize == 16
    assert vx.CORNER_DTYPE.itemsize == 12
    assert ctypes.sizeof(vx.Params) == 14 * 4
    assert vx.FRAME_INFO_DTYPE.itemsize == 40  # 3 doubles + 4 ints, no padding
    assert [vx.FRAME_INFO_DTYPE.fields[k][1] for k in ("mean", "stddev", "parallax", "grid_occupied", "grid_mask",
                                                       "n_keypoints", "n_matches")] == [0, 8, 16, 24, 28, 32, 36]


def test_create_rejects_bad_params():
    lib = vx.load_library()
    p = vx.Params()
    lib.vxorb_default_params(ctypes.byref(p))
    p.patch_size = 15
    h = ctypes.c_void_p()
    assert lib.vxorb_create(ctypes.byref(p), ctypes.byref(h)) == vx.E_INVALID
    assert not h.value
    assert b"patch_size" in lib.vxorb_last_error(None)


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_has_gpu(), reason="only meaningful on a box without a GPU")
def test_no_gpu_fails_loudly():
    """The product path must not silently fall back to CPU code."""
    with pytest.raises(vx.VxOrbError) as ei:
        vx.VxOrb()
    assert ei.value.code == vx.E_CUDA
    assert "no CPU fallback" in str(ei.value)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "visionx_slam_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".h", ".cuh", ".cpp", ".hpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text, f
                assert "orb_oracle" not in text and "liborb_oracle" not in text, f
