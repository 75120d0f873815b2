"""visionx_slam_b200 — B200-native ORB extraction + Hamming matching for VisionX-SLAM's front-end.

The product is the C-ABI shared library ``libvxorb.so`` (include/vxorb.h; CUDA sources in
``csrc/``).  This module is the thin Python host side used by the tests and the benchmark:

* :class:`VxOrb` — 1:1 binding of the C ABI (numpy arrays in, numpy arrays out).
* :class:`ORBExtractor`, :class:`ORBMatcher`, :class:`Frame`, :class:`Feature`, :class:`DMatch` —
  mirror of the reference's plug-in interface (core/feature/feature_extractor.h:10-16,
  feature_matcher.h:7-13, orb_extractor.h:9-20, orb_matcher.h:11-29, frame/frame.h:16-64) so the
  parity tests read like calls into the reference.  The C++ adapters a maintainer would compile
  into the reference live in ``cpp/``.

There is no CPU fallback: if ``libvxorb.so`` is missing or no sm_100 GPU is present, construction
raises.  Nothing here imports ``oracle``.
"""
import ctypes as C
import os

import numpy as np

from . import synth  # noqa: F401

_HERE = os.path.dirname(os.path.abspath(__file__))
# VXORB_LIB selects another build of the same library (A/B runs of compile-time variants, scratch/ab.sh)
LIB_PATH = os.environ.get("VXORB_LIB") or os.path.join(_HERE, "libvxorb.so")

KEYPOINT_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("response", "<f4"), ("angle", "<f4"),
                           ("octave", "<i4")])
MATCH_DTYPE = np.dtype([("query_idx", "<i4"), ("train_idx", "<i4"), ("distance", "<i4")])
KNN2_DTYPE = np.dtype([("idx", "<i4", (2,)), ("dist", "<i4", (2,))])
CORNER_DTYPE = np.dtype([("x", "<i4"), ("y", "<i4"), ("score", "<i4")])
FRAME_INFO_DTYPE = np.dtype([("mean", "<f8"), ("stddev", "<f8"), ("parallax", "<f8"), ("grid_occupied", "<i4"),
                             ("grid_mask", "<u4"), ("n_keypoints", "<i4"), ("n_matches", "<i4")])

OK, E_INVALID, E_CUDA, E_CAPACITY, E_NOMEM = 0, -1, -2, -3, -4
STAGES = ("pyramid", "fast", "select", "blur", "describe", "match")


class Params(C.Structure):
    _fields_ = [("n_features", C.c_int), ("scale_factor", C.c_float), ("n_levels", C.c_int),
                ("edge_threshold", C.c_int), ("patch_size", C.c_int), ("fast_threshold", C.c_int),
                ("nn_ratio", C.c_float), ("min_matches", C.c_int), ("device_id", C.c_int),
                ("max_width", C.c_int), ("max_height", C.c_int), ("max_batch", C.c_int),
                ("max_keypoints", C.c_int), ("max_train", C.c_int)]


class VxOrbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"vxorb error {code}: {msg}")
        self.code = code


# Every symbol include/vxorb.h declares: (name, restype, argtypes)
_vp, _i, _f, _sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
_SYMBOLS = [
    ("vxorb_default_params", None, [C.POINTER(Params)]),
    ("vxorb_create", _i, [C.POINTER(Params), C.POINTER(_vp)]),
    ("vxorb_destroy", None, [_vp]),
    ("vxorb_last_error", C.c_char_p, [_vp]),
    ("vxorb_version", C.c_char_p, []),
    ("vxorb_level_geometry", _i, [_vp, _i, _i, _vp, _vp, _vp, _vp]),
    ("vxorb_extract", _i, [_vp, _vp, _i, _i, _sz, _i, _vp, _vp, _vp]),
    ("vxorb_extract_bgr", _i, [_vp, _vp, _i, _i, _sz, _i, _vp, _vp, _vp]),
    ("vxorb_extract_info", _i, [_vp, _vp, _i, _i, _i, _sz, _i, _vp, _vp, _vp, _vp]),
    ("vxorb_extract_id", _i, [_vp, _vp, _i, _i, _i, _sz, _i, _vp, _vp, _vp, _vp, _vp]),
    ("vxorb_set_upload", _i, [_vp, _vp, _vp, _i, _vp]),
    ("vxorb_set_free", _i, [_vp, C.c_longlong]),
    ("vxorb_set_size", _i, [_vp, C.c_longlong]),
    ("vxorb_match_ids", _i, [_vp, C.c_longlong, C.c_longlong, _f, _f, _i, _i, _vp, _vp]),
    ("vxorb_extract_batch", _i, [_vp, _vp, _i, _i, _i, _sz, _i, _vp, _vp, _vp]),
    ("vxorb_knn2", _i, [_vp, _vp, _i, _vp, _i, _vp]),
    ("vxorb_match", _i, [_vp, _vp, _i, _vp, _i, _f, _i, _vp, _vp]),
    ("vxorb_knn2_windowed", _i, [_vp, _vp, _vp, _i, _vp, _vp, _i, _f, _vp]),
    ("vxorb_match_windowed", _i, [_vp, _vp, _vp, _i, _vp, _vp, _i, _f, _f, _i, _vp, _vp]),
    ("vxorb_knn2_boxed", _i, [_vp, _vp, _vp, _i, _vp, _vp, _i, _vp, _vp]),
    ("vxorb_match_boxed", _i, [_vp, _vp, _vp, _i, _vp, _vp, _i, _vp, _f, _i, _vp, _vp]),
    ("vxorb_knn2_masked", _i, [_vp, _vp, _i, _vp, _i, _vp, _sz, _vp]),
    ("vxorb_match_masked", _i, [_vp, _vp, _i, _vp, _i, _vp, _sz, _f, _i, _vp, _vp]),
    ("vxorb_knn2_static", _i, [_vp, _i, _vp, _i, _vp, _vp, _sz]),
    ("vxorb_depth_lookup", _i, [_vp, C.c_longlong, _vp, _i, _i, _i, _sz, _i, _vp, _vp]),
    ("vxorb_match_crosscheck", _i, [_vp, _vp, _i, _vp, _i, _i, _vp, _vp]),
    ("vxorb_match_filtered", _i, [_vp, _vp, _i, _vp, _i, _f, _i, _vp, _vp]),
    ("vxorb_track_slots", _i, [_vp]),
    ("vxorb_track_reset", _i, [_vp]),
    ("vxorb_track_submit", _i, [_vp, _vp, _i, _i, _i, _sz, _sz, _f, _i, _vp, _vp, _vp, _i, _vp, _vp, _vp]),
    ("vxorb_track_submit_info", _i, [_vp, _vp, _i, _i, _i, _sz, _sz, _f, _i, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp]),
    ("vxorb_track_submit_bgr", _i, [_vp, _vp, _i, _i, _i, _sz, _sz, _f, _i, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp]),
    ("vxorb_track_submit_stereo", _i, [_vp, _vp, _i, _i, _i, _sz, _sz, _vp, _f, _i, _vp, _vp, _vp, _i, _vp, _vp, _vp]),
    ("vxorb_track_wait", _i, [_vp, _i]),
    ("vxorb_keypoint_capacity", _i, [_vp]),
    ("vxorb_host_alloc", _vp, [_sz]),
    ("vxorb_host_free", None, [_vp]),
    ("vxorb_resident_upload", _i, [_vp, _vp, _i, _i, _i, _sz, _sz]),
    ("vxorb_resident_run", _i, [_vp, _f]),
    ("vxorb_resident_run_stereo", _i, [_vp, _vp, _f]),
    ("vxorb_resident_fetch", _i, [_vp, _i, _vp, _vp, _vp, _i, _vp, _vp]),
    ("vxorb_sync", _i, [_vp]),
    ("vxorb_timer_begin", _i, [_vp]),
    ("vxorb_timer_end", _i, [_vp, _vp]),
    ("vxorb_stage_times_reset", _i, [_vp, _i]),
    ("vxorb_stage_times", _i, [_vp, _vp, _vp]),
    ("vxorb_launch_count", C.c_longlong, [_vp]),
    ("vxorb_flush_l2", _i, [_vp]),
    ("vxorb_measure_popc_rate", _i, [_vp, _vp]),
    ("vxorb_debug_resize", _i, [_vp, _vp, _i, _i, _sz, _vp, _i, _i, _sz]),
    ("vxorb_debug_resize_band", _i, [_vp, _vp, _i, _i, _sz, _vp, _i, _i, _sz, _i]),
    ("vxorb_debug_fast", _i, [_vp, _vp, _i, _i, _sz, _i, _i, _i, _vp, _vp]),
    ("vxorb_debug_blur", _i, [_vp, _vp, _i, _i, _sz, _vp, _sz]),
    ("vxorb_debug_level", _i, [_vp, _i, _i, _i, _vp, _sz]),
]
EXPORTED_SYMBOLS = tuple(s[0] for s in _SYMBOLS)

_lib = None


def load_library(path=None):
    """dlopen libvxorb.so and bind every symbol of include/vxorb.h.  Raises if it is missing —
    build it with ``python -c 'import __graft_entry__ as g; g.build()'`` (or make -C csrc)."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise FileNotFoundError(f"{p} not found: the CUDA library is not built and there is no CPU fallback")
    lib = C.CDLL(p)
    for name, res, args in _SYMBOLS:
        fn = getattr(lib, name)  # AttributeError if the .so does not export what the header declares
        fn.restype = res
        fn.argtypes = args
    if path is None:
        _lib = lib
    return lib


def _ptr(a):
    return None if a is None else a.ctypes.data


def _desc(a):
    return np.ascontiguousarray(a, dtype=np.uint8).reshape(-1, 32)


def _xy(a):
    return np.ascontiguousarray(a, dtype=np.float32).reshape(-1, 2)


STATIC_MAX_QUERY, STATIC_MAX_TRAIN = 4096, 16384  # VXORB_STATIC_MAX_QUERY / _TRAIN of include/vxorb.h


def knn2_static(q, t):
    """vxorb_knn2_static: the rows of VxOrb.knn2 without a handle and without any allocation inside the call (the
    scratch is a static device array of the library).  Raises VxOrbError with the library's text beyond the limits."""
    lib = load_library()
    q, t = _desc(q), _desc(t)
    out = np.zeros(len(q), KNN2_DTYPE)
    err = C.create_string_buffer(256)
    rc = lib.vxorb_knn2_static(_ptr(q), len(q), _ptr(t), len(t), _ptr(out), C.addressof(err), len(err))
    if rc != 0:
        raise VxOrbError(rc, err.value.decode())
    return out


class PinnedArray:
    """numpy view of page-locked host memory from vxorb_host_alloc."""

    def __init__(self, shape, dtype):
        self._lib = load_library()
        dtype = np.dtype(dtype)
        n = int(np.prod(shape)) * dtype.itemsize
        self._p = self._lib.vxorb_host_alloc(max(n, 1))
        if not self._p:
            raise MemoryError("vxorb_host_alloc failed")
        buf = (C.c_uint8 * max(n, 1)).from_address(self._p)
        self.array = np.frombuffer(buf, dtype=np.uint8, count=n).view(dtype).reshape(shape)

    def free(self):
        if self._p:
            self.array = None
            self._lib.vxorb_host_free(self._p)
            self._p = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class VxOrb:
    """One handle of the C ABI (one GPU, one host thread)."""

    def __init__(self, n_features=1000, scale_factor=1.2, n_levels=8, fast_threshold=20, nn_ratio=0.8,
                 min_matches=50, device_id=0, max_width=640, max_height=480, max_batch=1, max_keypoints=0,
                 max_train=16384):
        self._lib = load_library()
        p = Params()
        self._lib.vxorb_default_params(C.byref(p))
        p.n_features, p.scale_factor, p.n_levels = n_features, scale_factor, n_levels
        p.fast_threshold, p.nn_ratio, p.min_matches = fast_threshold, nn_ratio, min_matches
        p.device_id, p.max_width, p.max_height = device_id, max_width, max_height
        p.max_batch, p.max_keypoints, p.max_train = max_batch, max_keypoints, max_train
        self.params = p
        self._h = _vp()
        rc = self._lib.vxorb_create(C.byref(p), C.byref(self._h))
        if rc != OK:
            msg = self._lib.vxorb_last_error(None).decode()
            self._h = None
            raise VxOrbError(rc, msg)
        self.kp_cap = self._lib.vxorb_keypoint_capacity(self._h)

    # -- plumbing --------------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None):
            self._lib.vxorb_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc, allow=()):
        if rc != OK and rc not in allow:
            raise VxOrbError(rc, self._lib.vxorb_last_error(self._h).decode())
        return rc

    def last_error(self):
        return self._lib.vxorb_last_error(self._h).decode()

    def level_geometry(self, w, h):
        n = self.params.n_levels
        lw, lh, q = (np.zeros(n, np.int32) for _ in range(3))
        ls = np.zeros(n, np.float32)
        self._check(self._lib.vxorb_level_geometry(self._h, w, h, _ptr(lw), _ptr(lh), _ptr(ls), _ptr(q)))
        return lw, lh, ls, q

    # -- extraction ------------------------------------------------------------------------------
    def extract(self, gray, kp_cap=None, return_code=False):
        """cv::ORB::detectAndCompute on one gray (H,W) or BGR (H,W,3) uint8 frame."""
        img = np.asarray(gray)
        if img.dtype != np.uint8 or img.ndim not in (2, 3):
            raise ValueError("expected uint8 (H,W) or (H,W,3)")
        if not img.flags.c_contiguous and not (img.ndim == 2 and img.strides[1] == 1 and img.strides[0] > 0):
            img = np.ascontiguousarray(img)
        cap = kp_cap if kp_cap is not None else self.kp_cap
        kp = np.zeros(cap, KEYPOINT_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = C.c_int(0)
        h, w = img.shape[:2]
        fn = self._lib.vxorb_extract_bgr if img.ndim == 3 else self._lib.vxorb_extract
        rc = fn(self._h, _ptr(img), w, h, img.strides[0], cap, _ptr(kp), _ptr(desc), C.byref(n))
        self._check(rc, allow=(E_CAPACITY,) if return_code else ())
        m = min(n.value, cap)
        out = (kp[:m].copy(), desc[:m].copy())
        return out + (rc, n.value) if return_code else out

    def extract_info(self, img):
        """extract() plus the frame's vxorb_frame_info: cv::meanStdDev of the gray image and the 5x5 grid
        coverage of its keypoints (Tracking::CheckImageQuality / CheckFeatureDistribution)."""
        img = np.ascontiguousarray(img, dtype=np.uint8)
        cap = self.kp_cap
        kp = np.zeros(cap, KEYPOINT_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        info = np.zeros(1, FRAME_INFO_DTYPE)
        n = C.c_int(0)
        h, w = img.shape[:2]
        self._check(self._lib.vxorb_extract_info(self._h, _ptr(img), 3 if img.ndim == 3 else 1, w, h, img.strides[0], cap,
                                                 _ptr(kp), _ptr(desc), C.byref(n), _ptr(info)))
        m = min(n.value, cap)
        return kp[:m].copy(), desc[:m].copy(), info[0].copy()

    def extract_into(self, img, kp, desc, want_id=False):
        """Raw single-frame call into caller-owned arrays (no allocation: what the latency figures time).
        Returns n, or (n, id) with the frame's descriptors kept resident on the device under id."""
        n = C.c_int(0)
        h, w = img.shape[:2]
        ch = 3 if img.ndim == 3 else 1
        if want_id:
            fid = C.c_longlong(0)
            self._check(self._lib.vxorb_extract_id(self._h, _ptr(img), ch, w, h, img.strides[0], len(kp), _ptr(kp), _ptr(desc),
                                                   C.byref(n), None, C.byref(fid)))
            return n.value, fid.value
        fn = self._lib.vxorb_extract_bgr if ch == 3 else self._lib.vxorb_extract
        self._check(fn(self._h, _ptr(img), w, h, img.strides[0], len(kp), _ptr(kp), _ptr(desc), C.byref(n)))
        return n.value

    def extract_id(self, img, with_info=False):
        """extract() that also returns the id under which the descriptors stay on the device."""
        img = np.asarray(img)
        if not img.flags.c_contiguous:
            img = np.ascontiguousarray(img)
        kp = np.zeros(self.kp_cap, KEYPOINT_DTYPE)
        desc = np.zeros((self.kp_cap, 32), np.uint8)
        info = np.zeros(1, FRAME_INFO_DTYPE)
        n, fid = C.c_int(0), C.c_longlong(0)
        h, w = img.shape[:2]
        self._check(self._lib.vxorb_extract_id(self._h, _ptr(img), 3 if img.ndim == 3 else 1, w, h, img.strides[0],
                                               self.kp_cap, _ptr(kp), _ptr(desc), C.byref(n),
                                               _ptr(info) if with_info else None, C.byref(fid)))
        m = min(n.value, self.kp_cap)
        out = (kp[:m].copy(), desc[:m].copy(), fid.value)
        return out + (info[0].copy(),) if with_info else out

    def set_upload(self, desc, xy=None):
        """Keeps descriptor rows (and optional positions) resident on the device; returns the set id."""
        desc = _desc(desc)
        xy = _xy(xy) if xy is not None else None
        sid = C.c_longlong(0)
        self._check(self._lib.vxorb_set_upload(self._h, _ptr(desc), _ptr(xy), len(desc), C.byref(sid)))
        return sid.value

    def set_free(self, sid):
        self._check(self._lib.vxorb_set_free(self._h, sid))

    def set_size(self, sid):
        return self._lib.vxorb_set_size(self._h, sid)

    def match_ids_into(self, id_q, id_t, out, radius=0.0, nn_ratio=0.0, mode=0):
        n = C.c_int(0)
        self._check(self._lib.vxorb_match_ids(self._h, id_q, id_t, radius, nn_ratio, mode, len(out), _ptr(out), C.byref(n)))
        return n.value

    def match_ids(self, id_q, id_t, radius=0.0, nn_ratio=0.0, mode=0):
        """vxorb_match / _filtered / _crosscheck (mode 0 / 1 / 2) on two device-resident sets."""
        nq = self.set_size(id_q)
        if nq < 0:
            raise VxOrbError(E_INVALID, "unknown or overwritten set id")
        out = np.zeros(max(nq, 1), MATCH_DTYPE)
        return out[:self.match_ids_into(id_q, id_t, out, radius, nn_ratio, mode)].copy()

    def match_into(self, q, t, out, nn_ratio=0.0):
        """Raw vxorb_match into a caller-owned array (q, t: contiguous (n,32) uint8)."""
        n = C.c_int(0)
        self._check(self._lib.vxorb_match(self._h, _ptr(q), len(q), _ptr(t), len(t), nn_ratio, len(out), _ptr(out), C.byref(n)))
        return n.value

    def extract_batch(self, frames, kp_cap=None):
        frames = [np.ascontiguousarray(f, dtype=np.uint8) for f in frames]
        b = len(frames)
        h, w = frames[0].shape
        cap = kp_cap if kp_cap is not None else self.kp_cap
        ptrs = (C.c_void_p * b)(*[f.ctypes.data for f in frames])
        kp = np.zeros((b, cap), KEYPOINT_DTYPE)
        desc = np.zeros((b, cap, 32), np.uint8)
        n = np.zeros(b, np.int32)
        self._check(self._lib.vxorb_extract_batch(self._h, C.cast(ptrs, _vp), b, w, h, w, cap, _ptr(kp), _ptr(desc), _ptr(n)))
        return [(kp[i, :n[i]].copy(), desc[i, :n[i]].copy()) for i in range(b)]

    # -- matching --------------------------------------------------------------------------------
    def knn2(self, q, t, q_xy=None, t_xy=None, radius=0.0):
        q, t = _desc(q), _desc(t)
        out = np.zeros(len(q), KNN2_DTYPE)
        if q_xy is not None:
            q_xy, t_xy = _xy(q_xy), _xy(t_xy)
            rc = self._lib.vxorb_knn2_windowed(self._h, _ptr(q), _ptr(q_xy), len(q), _ptr(t), _ptr(t_xy), len(t),
                                               radius, _ptr(out))
        else:
            rc = self._lib.vxorb_knn2(self._h, _ptr(q), len(q), _ptr(t), len(t), _ptr(out))
        self._check(rc)
        return out

    def _match_call(self, fn, q, nargs):
        out = np.zeros(max(len(q), 1), MATCH_DTYPE)
        n = C.c_int(0)
        self._check(fn(*nargs, len(out), _ptr(out), C.byref(n)))
        return out[:n.value].copy()

    def match(self, q, t, nn_ratio=0.0, q_xy=None, t_xy=None, radius=0.0):
        """ORBMatcher::Match semantics; windowed when positions and radius are given."""
        q, t = _desc(q), _desc(t)
        if q_xy is not None:
            q_xy, t_xy = _xy(q_xy), _xy(t_xy)
            return self._match_call(self._lib.vxorb_match_windowed, q,
                                    (self._h, _ptr(q), _ptr(q_xy), len(q), _ptr(t), _ptr(t_xy), len(t), radius, nn_ratio))
        return self._match_call(self._lib.vxorb_match, q, (self._h, _ptr(q), len(q), _ptr(t), len(t), nn_ratio))

    def knn2_boxed(self, q, t, q_xy, t_xy, box):
        q, t, q_xy, t_xy = _desc(q), _desc(t), _xy(q_xy), _xy(t_xy)
        b = np.ascontiguousarray(box, dtype=np.float32)
        out = np.zeros(len(q), KNN2_DTYPE)
        self._check(self._lib.vxorb_knn2_boxed(self._h, _ptr(q), _ptr(q_xy), len(q), _ptr(t), _ptr(t_xy), len(t), _ptr(b),
                                               _ptr(out)))
        return out

    def match_boxed(self, q, t, q_xy, t_xy, box, nn_ratio=0.0):
        """Stereo / epipolar-band matching: mask dx in [box0, box1], dy in [box2, box3] on (query - train)."""
        q, t, q_xy, t_xy = _desc(q), _desc(t), _xy(q_xy), _xy(t_xy)
        b = np.ascontiguousarray(box, dtype=np.float32)
        out = np.zeros(max(len(q), 1), MATCH_DTYPE)
        n = C.c_int(0)
        self._check(self._lib.vxorb_match_boxed(self._h, _ptr(q), _ptr(q_xy), len(q), _ptr(t), _ptr(t_xy), len(t), _ptr(b),
                                                nn_ratio, len(out), _ptr(out), C.byref(n)))
        return out[:n.value].copy()

    def knn2_masked(self, q, t, mask):
        """cv knnMatch(k=2, mask): mask (nq, nt) uint8, pair allowed iff non-zero."""
        q, t = _desc(q), _desc(t)
        mask = np.ascontiguousarray(mask, dtype=np.uint8).reshape(len(q), len(t))
        out = np.zeros(len(q), KNN2_DTYPE)
        self._check(self._lib.vxorb_knn2_masked(self._h, _ptr(q), len(q), _ptr(t), len(t), _ptr(mask), mask.strides[0], _ptr(out)))
        return out

    def match_masked(self, q, t, mask, nn_ratio=0.0):
        q, t = _desc(q), _desc(t)
        mask = np.ascontiguousarray(mask, dtype=np.uint8).reshape(len(q), len(t))
        out = np.zeros(max(len(q), 1), MATCH_DTYPE)
        n = C.c_int(0)
        self._check(self._lib.vxorb_match_masked(self._h, _ptr(q), len(q), _ptr(t), len(t), _ptr(mask), mask.strides[0],
                                                 nn_ratio, len(out), _ptr(out), C.byref(n)))
        return out[:n.value].copy()

    def depth_lookup(self, frame_id, depth):
        """Per-feature depth in metres of a resident frame (Tracking::CreateLandmarksFromDepth); 0 = skipped."""
        depth = np.ascontiguousarray(depth)
        dt = {np.dtype(np.uint16): 0, np.dtype(np.float32): 1, np.dtype(np.float64): 2}[depth.dtype]
        out = np.zeros(self.kp_cap, np.float64)
        n = C.c_int(0)
        self._check(self._lib.vxorb_depth_lookup(self._h, frame_id, _ptr(depth), dt, depth.shape[1], depth.shape[0],
                                                 depth.strides[0], len(out), _ptr(out), C.byref(n)))
        return out[:n.value].copy()

    def match_crosscheck(self, q, t):
        q, t = _desc(q), _desc(t)
        return self._match_call(self._lib.vxorb_match_crosscheck, q, (self._h, _ptr(q), len(q), _ptr(t), len(t)))

    def match_filtered(self, q, t, nn_ratio=0.0):
        q, t = _desc(q), _desc(t)
        return self._match_call(self._lib.vxorb_match_filtered, q, (self._h, _ptr(q), len(q), _ptr(t), len(t), nn_ratio))

    # -- streaming sequence path -----------------------------------------------------------------
    def track_reset(self):
        self._check(self._lib.vxorb_track_reset(self._h))

    def track_slots(self):
        return self._lib.vxorb_track_slots(self._h)

    def track_submit(self, frames, radius, kp_out=None, desc_out=None, n_kp_out=None, match_out=None,
                     n_match_out=None):
        """frames: (B,H,W) uint8, C-contiguous (pinned for overlap).  Output arrays (optional) must
        be (B,kp_cap) KEYPOINT, (B,kp_cap,32) u8, (B,) i32, (B,kp_cap) MATCH, (B,) i32."""
        b, h, w = frames.shape
        slot = C.c_int(-1)
        self._check(self._lib.vxorb_track_submit(self._h, _ptr(frames), b, w, h, frames.strides[1], frames.strides[0],
                                                 radius, self.kp_cap, _ptr(kp_out), _ptr(desc_out), _ptr(n_kp_out),
                                                 self.kp_cap, _ptr(match_out), _ptr(n_match_out), C.byref(slot)))
        return slot.value

    def track_wait(self, slot):
        self._check(self._lib.vxorb_track_wait(self._h, slot))

    def track_bgr(self, frames, radius):
        """track() for a (B,H,W,3) BGR chunk (converted on the device as cv::ORB converts a colour input)."""
        frames = np.asarray(frames)
        if frames.dtype != np.uint8 or frames.ndim != 4 or frames.strides[3] != 1 or frames.strides[2] != 3:
            frames = np.ascontiguousarray(frames, dtype=np.uint8)  # padded rows / frames are passed through as strides
        b, h, w, c = frames.shape
        assert c == 3
        kp = np.zeros((b, self.kp_cap), KEYPOINT_DTYPE)
        desc = np.zeros((b, self.kp_cap, 32), np.uint8)
        nk = np.zeros(b, np.int32)
        mt = np.zeros((b, self.kp_cap), MATCH_DTYPE)
        nm = np.zeros(b, np.int32)
        slot = C.c_int(-1)
        self._check(self._lib.vxorb_track_submit_bgr(self._h, _ptr(frames), b, w, h, frames.strides[1], frames.strides[0],
                                                     radius, self.kp_cap, _ptr(kp), _ptr(desc), _ptr(nk), self.kp_cap,
                                                     _ptr(mt), _ptr(nm), None, None, C.byref(slot)))
        self.track_wait(slot.value)
        return [(kp[i, :nk[i]].copy(), desc[i, :nk[i]].copy(), mt[i, :nm[i]].copy()) for i in range(b)]

    def track_info(self, frames, radius):
        """track() plus per-frame vxorb_frame_info and the matched positions (x_last, y_last, x_curr, y_curr)."""
        frames = np.ascontiguousarray(frames, dtype=np.uint8)
        b, h, w = frames.shape
        kp = np.zeros((b, self.kp_cap), KEYPOINT_DTYPE)
        desc = np.zeros((b, self.kp_cap, 32), np.uint8)
        nk = np.zeros(b, np.int32)
        mt = np.zeros((b, self.kp_cap), MATCH_DTYPE)
        nm = np.zeros(b, np.int32)
        info = np.zeros(b, FRAME_INFO_DTYPE)
        pts = np.zeros((b, self.kp_cap, 4), np.float32)
        slot = C.c_int(-1)
        self._check(self._lib.vxorb_track_submit_info(self._h, _ptr(frames), b, w, h, frames.strides[1], frames.strides[0],
                                                      radius, self.kp_cap, _ptr(kp), _ptr(desc), _ptr(nk), self.kp_cap,
                                                      _ptr(mt), _ptr(nm), _ptr(info), _ptr(pts), C.byref(slot)))
        self.track_wait(slot.value)
        return [(kp[i, :nk[i]].copy(), desc[i, :nk[i]].copy(), mt[i, :nm[i]].copy(), info[i].copy(), pts[i, :nm[i]].copy())
                for i in range(b)]

    def track(self, frames, radius):
        """Synchronous convenience: one chunk in, per-frame (kp, desc, matches) out."""
        frames = np.ascontiguousarray(frames, dtype=np.uint8)
        b = frames.shape[0]
        kp = np.zeros((b, self.kp_cap), KEYPOINT_DTYPE)
        desc = np.zeros((b, self.kp_cap, 32), np.uint8)
        nk = np.zeros(b, np.int32)
        mt = np.zeros((b, self.kp_cap), MATCH_DTYPE)
        nm = np.zeros(b, np.int32)
        slot = self.track_submit(frames, radius, kp, desc, nk, mt, nm)
        self.track_wait(slot)
        return [(kp[i, :nk[i]].copy(), desc[i, :nk[i]].copy(), mt[i, :nm[i]].copy()) for i in range(b)]

    def track_stereo(self, frames, box, radius):
        """One chunk of (left, right) pairs: frames (2P,H,W).  Returns per frame (kp, desc) and per pair
        (stereo matches left_k -> right_k, temporal matches left_(k-1) -> left_k)."""
        frames = np.ascontiguousarray(frames, dtype=np.uint8)
        b, h, w = frames.shape
        box = np.ascontiguousarray(box, dtype=np.float32)
        kp = np.zeros((b, self.kp_cap), KEYPOINT_DTYPE)
        desc = np.zeros((b, self.kp_cap, 32), np.uint8)
        nk = np.zeros(b, np.int32)
        mt = np.zeros((b, self.kp_cap), MATCH_DTYPE)
        nm = np.zeros(b, np.int32)
        slot = C.c_int(-1)
        self._check(self._lib.vxorb_track_submit_stereo(self._h, _ptr(frames), b // 2, w, h, frames.strides[1], frames.strides[0],
                                                        _ptr(box), radius, self.kp_cap, _ptr(kp), _ptr(desc), _ptr(nk),
                                                        self.kp_cap, _ptr(mt), _ptr(nm), C.byref(slot)))
        self.track_wait(slot.value)
        p = b // 2
        return ([(kp[i, :nk[i]].copy(), desc[i, :nk[i]].copy()) for i in range(b)],
                [(mt[k, :nm[k]].copy(), mt[p + k, :nm[p + k]].copy()) for k in range(p)])

    def resident_run_stereo(self, box, radius):
        box = np.ascontiguousarray(box, dtype=np.float32)
        self._check(self._lib.vxorb_resident_run_stereo(self._h, _ptr(box), radius))

    def resident_upload(self, frames):
        frames = np.ascontiguousarray(frames, dtype=np.uint8)
        b, h, w = frames.shape
        self._check(self._lib.vxorb_resident_upload(self._h, _ptr(frames), b, w, h, frames.strides[1], frames.strides[0]))
        self._resident_batch = b

    def resident_run(self, radius):
        self._check(self._lib.vxorb_resident_run(self._h, radius))

    def resident_fetch(self):
        b = self._resident_batch
        kp = np.zeros((b, self.kp_cap), KEYPOINT_DTYPE)
        desc = np.zeros((b, self.kp_cap, 32), np.uint8)
        nk = np.zeros(b, np.int32)
        mt = np.zeros((b, self.kp_cap), MATCH_DTYPE)
        nm = np.zeros(b, np.int32)
        self._check(self._lib.vxorb_resident_fetch(self._h, self.kp_cap, _ptr(kp), _ptr(desc), _ptr(nk), self.kp_cap,
                                                   _ptr(mt), _ptr(nm)))
        return [(kp[i, :nk[i]].copy(), desc[i, :nk[i]].copy(), mt[i, :nm[i]].copy()) for i in range(b)]

    # -- measurement -----------------------------------------------------------------------------
    def sync(self):
        self._check(self._lib.vxorb_sync(self._h))

    def timer_begin(self):
        self._check(self._lib.vxorb_timer_begin(self._h))

    def timer_end(self):
        ms = C.c_float(0)
        self._check(self._lib.vxorb_timer_end(self._h, C.byref(ms)))
        return ms.value

    def stage_times_reset(self, enable=True):
        self._check(self._lib.vxorb_stage_times_reset(self._h, 1 if enable else 0))

    def stage_times(self):
        ms = np.zeros(len(STAGES), np.float32)
        n = np.zeros(len(STAGES), np.int32)
        self._check(self._lib.vxorb_stage_times(self._h, _ptr(ms), _ptr(n)))
        return {s: (float(ms[i]), int(n[i])) for i, s in enumerate(STAGES)}

    def launch_count(self):
        return int(self._lib.vxorb_launch_count(self._h))

    def measure_popc_rate(self):
        """Sustained 32-bit popcount lane-ops per second of this GPU (register-only kernel)."""
        v = C.c_double(0)
        self._check(self._lib.vxorb_measure_popc_rate(self._h, C.byref(v)))
        return v.value

    def flush_l2(self):
        self._check(self._lib.vxorb_flush_l2(self._h))

    # -- stage isolation (parity tests) ----------------------------------------------------------
    def debug_resize(self, src, dw, dh):
        src = np.ascontiguousarray(src, dtype=np.uint8)
        sh, sw = src.shape
        dst = np.zeros((dh, dw), np.uint8)
        self._check(self._lib.vxorb_debug_resize(self._h, _ptr(src), sw, sh, sw, _ptr(dst), dw, dh, dw))
        return dst

    def debug_resize_band(self, src, dw, dh, band_rows):
        """One level through the production band kernel: band_rows 16 = throughput shape, 8 / 4 = latency shape."""
        src = np.ascontiguousarray(src, dtype=np.uint8)
        sh, sw = src.shape
        dst = np.zeros((dh, dw), np.uint8)
        self._check(self._lib.vxorb_debug_resize_band(self._h, _ptr(src), sw, sh, sw, _ptr(dst), dw, dh, dw, band_rows))
        return dst

    def debug_fast(self, img, threshold=20, border=3):
        img = np.ascontiguousarray(img, dtype=np.uint8)
        h, w = img.shape
        cap = (w // 2 + 2) * (h // 2 + 2)
        out = np.zeros(cap, CORNER_DTYPE)
        n = C.c_int(0)
        self._check(self._lib.vxorb_debug_fast(self._h, _ptr(img), w, h, w, threshold, border, cap, _ptr(out), C.byref(n)))
        return out[:n.value].copy()

    def debug_blur(self, img):
        img = np.ascontiguousarray(img, dtype=np.uint8)
        h, w = img.shape
        out = np.zeros_like(img)
        self._check(self._lib.vxorb_debug_blur(self._h, _ptr(img), w, h, w, _ptr(out), w))
        return out

    def debug_level(self, frame, level, w, h, blurred=False):
        """Pyramid (or blurred pyramid) level of frame `frame` of the last extract on w x h input."""
        lw, lh, _, _ = self.level_geometry(w, h)
        out = np.zeros((int(lh[level]), int(lw[level])), np.uint8)
        self._check(self._lib.vxorb_debug_level(self._h, frame, level, 1 if blurred else 0, _ptr(out), out.shape[1]))
        return out


# ---- mirror of the reference's plug-in interface -----------------------------------------------

class Feature:
    """core/frame/frame.h:16-23 — position (level-0 pixels, double) and response."""
    __slots__ = ("position", "response", "landmark_id_", "has_landmark", "is_outlier")

    def __init__(self, x, y, response):
        self.position = (float(x), float(y))
        self.response = float(response)
        self.landmark_id_ = 0
        self.has_landmark = False
        self.is_outlier = False


class Frame:
    """core/frame/frame.h:25-64 — the data carrier at the boundary (image, features, descriptors)."""

    def __init__(self, frame_id, timestamp, image):
        self._id, self._ts = frame_id, timestamp
        self._image = np.array(image, dtype=np.uint8, copy=True)  # frame.cpp:13 clones
        self._features = []
        self._descriptors = np.zeros((0, 32), np.uint8)

    def Id(self):
        return self._id

    def Image(self):
        return self._image

    def Features(self):
        return self._features

    def Descriptors(self):
        return self._descriptors


class DMatch:
    """cv::DMatch as ORBMatcher::Match fills it (orb_matcher.cpp:34)."""
    __slots__ = ("queryIdx", "trainIdx", "imgIdx", "distance")

    def __init__(self, q, t, d):
        self.queryIdx, self.trainIdx, self.imgIdx, self.distance = int(q), int(t), 0, float(d)


class ORBExtractor:
    """Drop-in for visionx::ORBExtractor (orb_extractor.h:9-20): same ctor arguments, Extract(frame)
    replaces frame.Features() and frame.Descriptors() (orb_extractor.cpp:9-27)."""

    def __init__(self, n_features=1000, scale_factor=1.2, n_levels=8, **handle_kw):
        handle_kw.setdefault("max_width", 1920)
        handle_kw.setdefault("max_height", 1088)
        self._kw = dict(handle_kw, n_features=n_features, scale_factor=scale_factor, n_levels=n_levels)
        self._vx = VxOrb(**self._kw)
        self.rebuilds = 0  # handles rebuilt for frames beyond max_width x max_height (cv::ORB takes any size)

    def _reserve(self, width, height):
        if width <= self._kw["max_width"] and height <= self._kw["max_height"]:
            return
        kw = dict(self._kw, max_width=max(self._kw["max_width"], (width + 63) // 64 * 64),
                  max_height=max(self._kw["max_height"], (height + 63) // 64 * 64))
        bigger = VxOrb(**kw)  # raises, and keeps the old handle, if the device cannot hold it
        self._vx.close()
        self._vx, self._kw = bigger, kw
        self.rebuilds += 1  # ids handed out by the old handle are void from here on

    def Extract(self, frame):
        img = frame.Image()
        self._reserve(img.shape[1], img.shape[0])
        kp, desc, fid = self._vx.extract_id(img)
        feats = frame.Features()
        feats.clear()
        feats.extend(Feature(k["x"], k["y"], k["response"]) for k in kp)
        frame._descriptors = desc
        frame._keypoints = kp
        # the descriptors also stay on the device under this id (ring of 8 frames) of this generation of the handle
        frame._vx_id = (self.rebuilds, fid)

    def resident_id(self, frame):
        """Device-resident id of the frame's descriptors, or 0 when they are no longer (or never were) there."""
        gen, fid = getattr(frame, "_vx_id", None) or (-1, 0)
        return fid if fid and gen == self.rebuilds and self._vx.set_size(fid) == len(frame.Descriptors()) else 0


class ORBMatcher:
    """Drop-in for visionx::ORBMatcher (orb_matcher.h:11-29): Options{nn_ratio=0.8, min_matches=50},
    Match(last, curr, matches) -> count, matches in ascending queryIdx order."""

    class Options:
        def __init__(self, nn_ratio=0.8, min_matches=50):
            self.nn_ratio, self.min_matches = nn_ratio, min_matches

    def __init__(self, options=None, extractor=None, **handle_kw):
        """extractor: an ORBExtractor whose handle is shared — frames it extracted are matched from the descriptors
        it left on the device (vxorb_match_ids); other descriptor arrays are uploaded."""
        self.options_ = options or ORBMatcher.Options()
        self._extractor = extractor
        self._own = None if extractor is not None else VxOrb(
            nn_ratio=self.options_.nn_ratio, min_matches=self.options_.min_matches, **handle_kw)
        self.warnings = []
        self.resident_calls = self.upload_calls = 0

    @property
    def _vx(self):
        # read per call: the extractor rebuilds its handle when a larger frame arrives
        return self._extractor._vx if self._extractor is not None else self._own

    def Match(self, last, curr, matches):
        matches.clear()
        d1, d2 = last.Descriptors(), curr.Descriptors()
        if len(d1) == 0 or len(d2) == 0:  # orb_matcher.cpp:18-20
            return 0
        i1 = self._extractor.resident_id(last) if self._extractor else 0
        i2 = self._extractor.resident_id(curr) if self._extractor else 0
        if i1 and i2:
            m = self._vx.match_ids(i1, i2, 0.0, self.options_.nn_ratio)
            self.resident_calls += 1
        else:
            m = self._vx.match(d1, d2, self.options_.nn_ratio)
            self.upload_calls += 1
        matches.extend(DMatch(r["query_idx"], r["train_idx"], r["distance"]) for r in m)
        if len(matches) < self.options_.min_matches:  # orb_matcher.cpp:38-40 (warning only)
            self.warnings.append(f"[ORBMatcher] Too few matches: {len(matches)}")
        return len(matches)
