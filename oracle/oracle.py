"""ctypes binding of oracle/liborb_oracle.so (TEST INFRASTRUCTURE ONLY).

Each function cites the reference behaviour it restates via orb_oracle.c; the reference call
sites are core/feature/orb_extractor.cpp:13 (cv::ORB::detectAndCompute) and
core/feature/orb_matcher.cpp:22-36 (cv::BFMatcher knnMatch + ratio test).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liborb_oracle.so")

KEYPOINT_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("response", "<f4"), ("angle", "<f4"),
                           ("octave", "<i4")])
MATCH_DTYPE = np.dtype([("query_idx", "<i4"), ("train_idx", "<i4"), ("distance", "<i4")])
KNN2_DTYPE = np.dtype([("idx", "<i4", (2,)), ("dist", "<i4", (2,))])
CORNER_DTYPE = np.dtype([("x", "<i4"), ("y", "<i4"), ("score", "<i4")])


def build(force=False):
    """Compile the C oracle with its Makefile (needs gcc; a no-op when the .so is current)."""
    src = os.path.join(_HERE, "orb_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        u8p, i32p, f32p, vp = C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p
        sz = C.c_size_t
        L.vxo_bgr2gray.argtypes = [u8p, C.c_int, C.c_int, sz, u8p, sz]
        L.vxo_bgr2gray.restype = None
        L.vxo_level_geometry.argtypes = [C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, i32p, i32p, f32p, i32p]
        L.vxo_level_geometry.restype = C.c_int
        L.vxo_resize_linear_exact.argtypes = [u8p, C.c_int, C.c_int, sz, u8p, C.c_int, C.c_int, sz]
        L.vxo_resize_linear_exact.restype = None
        L.vxo_fast9_nms.argtypes = [u8p, C.c_int, C.c_int, sz, C.c_int, C.c_int, vp, C.c_int]
        L.vxo_fast9_nms.restype = C.c_int
        L.vxo_harris.argtypes = [u8p, sz, C.c_int, C.c_int]
        L.vxo_harris.restype = C.c_float
        L.vxo_fast_atan2.argtypes = [C.c_float, C.c_float]
        L.vxo_fast_atan2.restype = C.c_float
        L.vxo_ic_angle.argtypes = [u8p, sz, C.c_int, C.c_int]
        L.vxo_ic_angle.restype = C.c_float
        L.vxo_gauss_kernel7.argtypes = [f32p]
        L.vxo_gauss_kernel7.restype = None
        L.vxo_blur7.argtypes = [u8p, C.c_int, C.c_int, sz, u8p, sz]
        L.vxo_blur7.restype = None
        L.vxo_brief.argtypes = [u8p, sz, C.c_int, C.c_int, C.c_float, u8p]
        L.vxo_brief.restype = None
        L.vxo_orb_extract.argtypes = [u8p, C.c_int, C.c_int, sz, C.c_int, C.c_float, C.c_int, C.c_int, vp, u8p, C.c_int]
        L.vxo_orb_extract.restype = C.c_int
        L.vxo_pyramid_level.argtypes = [u8p, C.c_int, C.c_int, sz, C.c_float, C.c_int, C.c_int, u8p]
        L.vxo_pyramid_level.restype = C.c_int
        L.vxo_knn2.argtypes = [u8p, f32p, C.c_int, u8p, f32p, C.c_int, C.c_float, vp]
        L.vxo_knn2.restype = None
        L.vxo_match.argtypes = [u8p, f32p, C.c_int, u8p, f32p, C.c_int, C.c_float, C.c_float, vp, C.c_int]
        L.vxo_match.restype = C.c_int
        L.vxo_match_crosscheck.argtypes = [u8p, C.c_int, u8p, C.c_int, vp, C.c_int]
        L.vxo_match_crosscheck.restype = C.c_int
        L.vxo_postfilter.argtypes = [vp, C.c_int, vp]
        L.vxo_postfilter.restype = C.c_int
        L.vxo_knn2_boxed.argtypes = [u8p, f32p, C.c_int, u8p, f32p, C.c_int, f32p, vp]
        L.vxo_knn2_boxed.restype = None
        L.vxo_match_boxed.argtypes = [u8p, f32p, C.c_int, u8p, f32p, C.c_int, f32p, C.c_float, vp, C.c_int]
        L.vxo_match_boxed.restype = C.c_int
        L.vxo_knn2_masked.argtypes = [u8p, C.c_int, u8p, C.c_int, u8p, sz, vp]
        L.vxo_knn2_masked.restype = None
        L.vxo_depth_lookup.argtypes = [f32p, C.c_int, vp, C.c_int, C.c_int, C.c_int, sz, vp]
        L.vxo_depth_lookup.restype = None
        L.vxo_image_stats.argtypes = [u8p, C.c_int, C.c_int, sz, vp, vp]
        L.vxo_image_stats.restype = None
        L.vxo_grid_coverage.argtypes = [f32p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp]
        L.vxo_grid_coverage.restype = C.c_int
        L.vxo_correspondences.argtypes = [f32p, f32p, vp, C.c_int, f32p]
        L.vxo_correspondences.restype = C.c_double
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data if a is not None else None


def _img(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    assert a.ndim == 2
    return a


def bgr2gray(bgr):
    bgr = np.ascontiguousarray(bgr, dtype=np.uint8)
    h, w, _ = bgr.shape
    out = np.empty((h, w), np.uint8)
    lib().vxo_bgr2gray(_p(bgr), w, h, w * 3, _p(out), w)
    return out


def level_geometry(W, H, n_features=1000, scale_factor=1.2, n_levels=8):
    lw = np.zeros(n_levels, np.int32)
    lh = np.zeros(n_levels, np.int32)
    ls = np.zeros(n_levels, np.float32)
    q = np.zeros(n_levels, np.int32)
    rc = lib().vxo_level_geometry(W, H, n_features, scale_factor, n_levels, _p(lw), _p(lh), _p(ls), _p(q))
    assert rc == 0
    return lw, lh, ls, q


def resize_linear_exact(src, dw, dh):
    src = _img(src)
    sh, sw = src.shape
    dst = np.empty((dh, dw), np.uint8)
    lib().vxo_resize_linear_exact(_p(src), sw, sh, sw, _p(dst), dw, dh, dw)
    return dst


def pyramid(gray, scale_factor=1.2, n_levels=8):
    gray = _img(gray)
    H, W = gray.shape
    lw, lh, _, _ = level_geometry(W, H, 0, scale_factor, n_levels)
    levels = [gray.copy()]
    for l in range(1, n_levels):
        levels.append(resize_linear_exact(levels[-1], int(lw[l]), int(lh[l])))
    return levels


def fast9_nms(img, threshold=20, border=3):
    img = _img(img)
    h, w = img.shape
    cap = (w // 2 + 1) * (h // 2 + 1)
    out = np.zeros(cap, CORNER_DTYPE)
    n = lib().vxo_fast9_nms(_p(img), w, h, w, threshold, border, _p(out), cap)
    assert n <= cap
    return out[:n].copy()


def harris(img, x, y):
    img = _img(img)
    return np.float32(lib().vxo_harris(_p(img), img.shape[1], int(x), int(y)))


def ic_angle(img, x, y):
    img = _img(img)
    return np.float32(lib().vxo_ic_angle(_p(img), img.shape[1], int(x), int(y)))


def fast_atan2(y, x):
    return np.float32(lib().vxo_fast_atan2(float(np.float32(y)), float(np.float32(x))))


def gauss_kernel7():
    k = np.zeros(7, np.float32)
    lib().vxo_gauss_kernel7(_p(k))
    return k


def blur7(img):
    img = _img(img)
    h, w = img.shape
    out = np.empty_like(img)
    lib().vxo_blur7(_p(img), w, h, w, _p(out), w)
    return out


def brief(blurred, cx, cy, angle_deg):
    blurred = _img(blurred)
    d = np.zeros(32, np.uint8)
    lib().vxo_brief(_p(blurred), blurred.shape[1], int(cx), int(cy), float(np.float32(angle_deg)), _p(d))
    return d


def orb_extract(gray, n_features=1000, scale_factor=1.2, n_levels=8, fast_threshold=20, cap=None):
    """cv::ORB(n_features, scale, levels).detectAndCompute in canonical (octave, y, x) order."""
    gray = _img(gray)
    H, W = gray.shape
    if cap is None:
        cap = max(4 * n_features, 4096)
    while True:
        kp = np.zeros(cap, KEYPOINT_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = lib().vxo_orb_extract(_p(gray), W, H, W, n_features, scale_factor, n_levels, fast_threshold,
                                  _p(kp), _p(desc), cap)
        assert n >= 0
        if n <= cap:
            return kp[:n].copy(), desc[:n].copy()
        cap = n


def _desc(a):
    a = np.ascontiguousarray(a, dtype=np.uint8).reshape(-1, 32)
    return a


def _xy(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32).reshape(-1, 2)


def knn2(q, t, q_xy=None, t_xy=None, radius=0.0):
    q, t = _desc(q), _desc(t)
    q_xy, t_xy = _xy(q_xy), _xy(t_xy)
    out = np.zeros(len(q), KNN2_DTYPE)
    lib().vxo_knn2(_p(q), _p(q_xy), len(q), _p(t), _p(t_xy), len(t), radius, _p(out))
    return out


def match(q, t, nn_ratio=0.8, q_xy=None, t_xy=None, radius=0.0):
    """ORBMatcher::Match (orb_matcher.cpp:11-43), optionally with the square window mask."""
    q, t = _desc(q), _desc(t)
    q_xy, t_xy = _xy(q_xy), _xy(t_xy)
    out = np.zeros(max(len(q), 1), MATCH_DTYPE)
    n = lib().vxo_match(_p(q), _p(q_xy), len(q), _p(t), _p(t_xy), len(t), radius, nn_ratio, _p(out), len(out))
    return out[:n].copy()


def match_crosscheck(q, t):
    q, t = _desc(q), _desc(t)
    out = np.zeros(max(len(q), 1), MATCH_DTYPE)
    n = lib().vxo_match_crosscheck(_p(q), len(q), _p(t), len(t), _p(out), len(out))
    return out[:n].copy()


def postfilter(matches):
    m = np.ascontiguousarray(matches, dtype=MATCH_DTYPE)
    out = np.zeros(max(len(m), 1), MATCH_DTYPE)
    n = lib().vxo_postfilter(_p(m), len(m), _p(out))
    return out[:n].copy()


def image_stats(gray):
    """cv::meanStdDev of a gray image (Tracking::CheckImageQuality, tracking.cpp:120-139) -> (mean, stddev)."""
    g = _img(gray)
    out = np.zeros(2, np.float64)
    lib().vxo_image_stats(_p(g), g.shape[1], g.shape[0], g.strides[0], out.ctypes.data, out.ctypes.data + 8)
    return float(out[0]), float(out[1])


def grid_coverage(xy, width, height, cols=5, rows=5):
    """Tracking::CheckFeatureDistribution (tracking.cpp:93-118) -> (occupied cells, mask)."""
    xy = _xy(xy) if xy is not None and len(xy) else np.zeros((0, 2), np.float32)
    mask = np.zeros(1, np.uint32)
    n = lib().vxo_grid_coverage(_p(xy), len(xy), width, height, cols, rows, _p(mask))
    return int(n), int(mask[0])


def correspondences(q_xy, t_xy, matches):
    """Matched point pairs (tracking.cpp:505-514) and ComputeParallax (tracking.cpp:548-560) -> (pts[n,4], parallax)."""
    m = np.ascontiguousarray(matches, dtype=MATCH_DTYPE)
    q_xy, t_xy = _xy(q_xy), _xy(t_xy)
    pts = np.zeros((max(len(m), 1), 4), np.float32)
    par = lib().vxo_correspondences(_p(q_xy), _p(t_xy), _p(m), len(m), _p(pts))
    return pts[:len(m)].copy(), float(par)


def knn2_masked(q, t, mask):
    """cv knnMatch(k=2, mask): pair (i, j) allowed iff mask[i, j] != 0."""
    q, t = _desc(q), _desc(t)
    mask = np.ascontiguousarray(mask, dtype=np.uint8).reshape(len(q), len(t))
    out = np.zeros(len(q), KNN2_DTYPE)
    lib().vxo_knn2_masked(_p(q), len(q), _p(t), len(t), _p(mask), mask.shape[1], _p(out))
    return out


def match_from_knn(knn, nn_ratio=0.8):
    """The ratio loop of ORBMatcher::Match (orb_matcher.cpp:27-36) over a 2-NN table."""
    r = np.float32(nn_ratio)
    rows = [(i, int(k["idx"][0]), int(k["dist"][0])) for i, k in enumerate(knn)
            if k["idx"][1] >= 0 and np.float32(k["dist"][0]) < r * np.float32(k["dist"][1])]
    return np.array(rows, MATCH_DTYPE) if rows else np.zeros(0, MATCH_DTYPE)


DEPTH_TYPES = {np.dtype(np.uint16): 0, np.dtype(np.float32): 1, np.dtype(np.float64): 2}


def depth_lookup(xy, depth):
    """Tracking::CreateLandmarksFromDepth's per-feature depth in metres (tracking.cpp:606-637); 0 = skipped."""
    xy = _xy(xy) if xy is not None and len(xy) else np.zeros((0, 2), np.float32)
    depth = np.ascontiguousarray(depth)
    out = np.zeros(max(len(xy), 1), np.float64)
    lib().vxo_depth_lookup(_p(xy), len(xy), depth.ctypes.data, DEPTH_TYPES[depth.dtype], depth.shape[1], depth.shape[0],
                           depth.strides[0], _p(out))
    return out[:len(xy)].copy()


def knn2_boxed(q, t, q_xy, t_xy, box):
    """cv knnMatch(k=2) with the box mask dx in [box0, box1], dy in [box2, box3] on (query - train)."""
    q, t, q_xy, t_xy = _desc(q), _desc(t), _xy(q_xy), _xy(t_xy)
    b = np.ascontiguousarray(box, dtype=np.float32)
    out = np.zeros(len(q), KNN2_DTYPE)
    lib().vxo_knn2_boxed(_p(q), _p(q_xy), len(q), _p(t), _p(t_xy), len(t), _p(b), _p(out))
    return out


def match_boxed(q, t, q_xy, t_xy, box, nn_ratio=0.8):
    q, t, q_xy, t_xy = _desc(q), _desc(t), _xy(q_xy), _xy(t_xy)
    b = np.ascontiguousarray(box, dtype=np.float32)
    out = np.zeros(max(len(q), 1), MATCH_DTYPE)
    n = lib().vxo_match_boxed(_p(q), _p(q_xy), len(q), _p(t), _p(t_xy), len(t), _p(b), nn_ratio, _p(out), len(out))
    return out[:n].copy()
