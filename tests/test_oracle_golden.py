"""The CPU oracle against the cv2-generated golden vectors (tests/golden/make_golden.py).

This is what pins the oracle: the reference has no tests of its own for this path, so the pin is
the output of the OpenCV entry points the reference calls (orb_extractor.cpp:13,
orb_matcher.cpp:22-36).  Runs on CPU.
"""
import numpy as np
import pytest

import oracle
from conftest import assert_kp_equal, load_golden, matches_as_rows
from visionx_slam_b200.synth import sha256, synth_scene


def test_level_geometry_and_quotas():
    lw, lh, ls, q = oracle.level_geometry(640, 480, 1000)
    assert list(lw) == [640, 533, 444, 370, 309, 257, 214, 179]
    assert list(lh) == [480, 400, 333, 278, 231, 193, 161, 134]
    assert list(q) == [217, 181, 151, 126, 105, 87, 73, 60]
    lw, lh, _, q = oracle.level_geometry(1241, 376, 2000)
    assert list(lw) == [1241, 1034, 862, 718, 598, 499, 416, 346]
    assert list(q) == [434, 362, 302, 251, 209, 175, 145, 122]
    lw, lh, _, q = oracle.level_geometry(1920, 1080, 4000)
    assert list(lh) == [1080, 900, 750, 625, 521, 434, 362, 301]
    assert list(q) == [869, 724, 603, 503, 419, 349, 291, 242]


def test_stage_pyramid_fast_blur_c1(golden_c1):
    g = golden_c1
    levels = oracle.pyramid(g["frame_a"])
    for l, lvl in enumerate(levels):
        assert tuple(g[f"a_level{l}_shape"]) == lvl.shape
        assert str(g[f"a_level{l}_sha"]) == sha256(lvl), f"pyramid level {l}"
        assert str(g[f"a_blur{l}_sha"]) == sha256(oracle.blur7(lvl)), f"blur level {l}"
        c = oracle.fast9_nms(lvl, 20, 3)
        ref = g[f"a_fast{l}"]
        got = np.stack([c["x"], c["y"], c["score"]], 1).astype(np.int16)
        assert np.array_equal(got, ref), f"FAST level {l}"


def test_extract_c1(golden_c1):
    g = golden_c1
    for tag in "ab":
        kp, desc = oracle.orb_extract(g["frame_" + tag])
        assert_kp_equal(kp, g["kp_" + tag])
        assert np.array_equal(desc, g["desc_" + tag])


@pytest.mark.parametrize("name", ["c3_1241x376.npz", "c4_1920x1080.npz"])
def test_extract_large_shapes(name):
    g = load_golden(name)
    img = synth_scene(int(g["w"]), int(g["h"]), int(g["seed"]))
    if sha256(img) != str(g["frame_sha"]):
        pytest.skip("synthetic generator not byte-reproducible on this numpy build")
    kp, desc = oracle.orb_extract(img, int(g["n_features"]))
    assert_kp_equal(kp, g["kp"])
    assert np.array_equal(desc, g["desc"])
    levels = oracle.pyramid(img)
    for l in (0, 7):
        c = oracle.fast9_nms(levels[l], 20, 3)
        got = np.stack([c["x"], c["y"], c["score"]], 1).astype(np.int16)
        assert np.array_equal(got, g[f"fast{l}"])
    for l in range(8):
        assert str(g[f"level{l}_sha"]) == sha256(levels[l])
        assert str(g[f"blur{l}_sha"]) == sha256(oracle.blur7(levels[l]))


@pytest.mark.parametrize("name", ["grid", "noise", "blank", "tiny60", "small80x70", "lowtex", "odd_357x291"])
def test_extract_adversarial(golden_adv, name):
    g = golden_adv
    kp, desc = oracle.orb_extract(g[name + "_img"])
    assert_kp_equal(kp, g[name + "_kp"])
    assert np.array_equal(desc, g[name + "_desc"])


@pytest.mark.parametrize("tag", ["p500_4", "p300_s15_5", "p2000_8"])
def test_extract_other_params(golden_adv, tag):
    g = golden_adv
    n, sf, nl = g[tag + "_params"]
    img = synth_scene(640, 480, int(g["params_img_seed"]))
    kp, desc = oracle.orb_extract(img, int(n), float(sf), int(nl))
    assert_kp_equal(kp, g[tag + "_kp"])
    assert np.array_equal(desc, g[tag + "_desc"])


def test_bgr_gray(golden_adv):
    g = golden_adv
    gray = oracle.bgr2gray(g["bgr_img"])
    assert np.array_equal(gray, g["bgr_gray"])
    kp, desc = oracle.orb_extract(gray, 500)
    assert_kp_equal(kp, g["bgr_kp"])
    assert np.array_equal(desc, g["bgr_desc"])


def _knn_equal(a, b):
    return np.array_equal(a["idx"], b["idx"]) and np.array_equal(
        np.where(a["idx"] >= 0, a["dist"], 0), np.where(b["idx"] >= 0, b["dist"], 0))


def test_match_c1(golden_c1):
    g = golden_c1
    knn = oracle.knn2(g["desc_a"], g["desc_b"])
    assert _knn_equal(knn, g["knn"])
    assert np.array_equal(matches_as_rows(oracle.match(g["desc_a"], g["desc_b"])), g["matches"])
    qxy = np.stack([g["kp_a"]["x"], g["kp_a"]["y"]], 1)
    txy = np.stack([g["kp_b"]["x"], g["kp_b"]["y"]], 1)
    assert _knn_equal(oracle.knn2(g["desc_a"], g["desc_b"], qxy, txy, 32.0), g["knn_win32"])
    m = oracle.match(g["desc_a"], g["desc_b"], 0.8, qxy, txy, 32.0)
    assert np.array_equal(matches_as_rows(m), g["matches_win32"])
    assert np.array_equal(matches_as_rows(oracle.match_crosscheck(g["desc_a"], g["desc_b"])), g["crosscheck"])


@pytest.mark.parametrize("tag", ["ties", "nt1", "nt2", "rand", "dup"])
def test_match_kat(golden_kat, tag):
    g = golden_kat
    q, t = g[tag + "_q"], g[tag + "_t"]
    assert _knn_equal(oracle.knn2(q, t), g[tag + "_knn"])
    assert np.array_equal(matches_as_rows(oracle.match(q, t)), g[tag + "_matches"])
    assert np.array_equal(matches_as_rows(oracle.match_crosscheck(q, t)), g[tag + "_crosscheck"])


def test_match_windowed_kat(golden_kat):
    g = golden_kat
    knn = oracle.knn2(g["dup_q"], g["dup_t"], g["dup_qxy"], g["dup_txy"], 64.0)
    assert _knn_equal(knn, g["dup_knn_win64"])
    m = oracle.match(g["dup_q"], g["dup_t"], 0.8, g["dup_qxy"], g["dup_txy"], 64.0)
    assert np.array_equal(matches_as_rows(m), g["dup_matches_win64"])


def test_match_empty():
    z = np.zeros((0, 32), np.uint8)
    q = np.zeros((5, 32), np.uint8)
    assert len(oracle.match(z, q)) == 0 and len(oracle.match(q, z)) == 0


def test_ratio_integer_equivalence():
    """d1 < 0.8f*d2 in float == 5*d1 < 4*d2 in integers for every 0 <= d <= 256."""
    d = np.arange(257)
    d1, d2 = np.meshgrid(d, d, indexing="ij")
    f = d1.astype(np.float32) < np.float32(0.8) * d2.astype(np.float32)
    assert np.array_equal(f, 5 * d1 < 4 * d2)


def test_postfilter():
    m = np.zeros(5, oracle.MATCH_DTYPE)
    m["distance"] = [10, 20, 31, 40, 25]
    m["query_idx"] = np.arange(5)
    out = oracle.postfilter(m)  # min 10 -> limit max(20, 30) = 30
    assert list(out["query_idx"]) == [0, 1, 4]
    m["distance"] = [40, 90, 79, 81, 80]
    assert list(oracle.postfilter(m)["query_idx"]) == [0, 2, 4]


# ---- caller-side per-frame quantities ----------------------------------------

def test_image_stats_golden():
    """Oracle restatement of cv::meanStdDev (Tracking::CheckImageQuality) is bit-exact against cv2."""
    g = load_golden("caller_kat.npz")
    for k in ("synth", "noise_odd", "const", "bright", "dark_w3", "bgr"):
        img = g[k + "_img"]
        gray = oracle.bgr2gray(img) if img.ndim == 3 else img
        mean, sd = oracle.image_stats(gray)
        assert mean == g[k + "_stats"][0] and sd == g[k + "_stats"][1], k


def test_grid_coverage_restatement():
    """Oracle vs a literal Python transcription of CheckFeatureDistribution (tracking.cpp:93-118)."""
    rng = np.random.default_rng(4)
    for n, (w, h) in [(0, (640, 480)), (1, (640, 480)), (40, (640, 480)), (1000, (1241, 376)), (7, (357, 291))]:
        xy = (rng.uniform(0, 1, (n, 2)) * np.array([w, h]) * rng.choice([0.2, 1.0])).astype(np.float32)
        if n > 2:
            xy[0] = (w, h)           # past the last cell: clamped
            xy[1] = (w * 0.6, 0.0)   # exactly on a cell boundary
        grid = [[False] * 5 for _ in range(5)]
        for x, y in xy:
            col = min(max(int((float(x) / w) * 5), 0), 4)
            row = min(max(int((float(y) / h) * 5), 0), 4)
            grid[col][row] = True
        want = sum(sum(r) for r in grid)
        got, mask = oracle.grid_coverage(xy, w, h)
        assert got == want and bin(mask).count("1") == want
        assert all(bool(mask >> (c * 5 + r) & 1) == grid[c][r] for c in range(5) for r in range(5))


def test_correspondences_restatement():
    rng = np.random.default_rng(9)
    qxy = rng.uniform(0, 640, (300, 2)).astype(np.float32)
    txy = rng.uniform(0, 640, (280, 2)).astype(np.float32)
    m = np.zeros(120, oracle.MATCH_DTYPE)
    m["query_idx"] = np.sort(rng.permutation(300)[:120])
    m["train_idx"] = rng.integers(0, 280, 120)
    pts, par = oracle.correspondences(qxy, txy, m)
    assert np.array_equal(pts[:, :2], qxy[m["query_idx"]]) and np.array_equal(pts[:, 2:], txy[m["train_idx"]])
    d = qxy[m["query_idx"]].astype(np.float64) - txy[m["train_idx"]].astype(np.float64)
    total = 0.0
    for dx, dy in d:  # serial sum in match order, as ComputeParallax (tracking.cpp:548-560)
        total += float(np.sqrt(dx * dx + dy * dy))
    assert par == total / 120
    assert oracle.correspondences(qxy, txy, m[:0])[1] == 0.0


# ---- BASELINE config 3: stereo left -> right matching with the epipolar-band box mask ---------------

def test_stereo_boxed_golden():
    """Oracle extract of both views + box-masked 2-NN + ratio against cv2 (stereo_kat.npz)."""
    from visionx_slam_b200.synth import synth_stereo_pair
    g = load_golden("stereo_kat.npz")
    left, right = synth_stereo_pair(1241, 376, 31)
    assert sha256(left) == str(g["left_sha"]) and sha256(right) == str(g["right_sha"])
    kl, dl = oracle.orb_extract(left, n_features=2000)
    kr, dr = oracle.orb_extract(right, n_features=2000)
    assert_kp_equal(kl, g["kp_left"])
    assert_kp_equal(kr, g["kp_right"])
    assert np.array_equal(dl, g["desc_left"]) and np.array_equal(dr, g["desc_right"])
    qxy, txy = np.stack([kl["x"], kl["y"]], 1), np.stack([kr["x"], kr["y"]], 1)
    knn = oracle.knn2_boxed(dl, dr, qxy, txy, g["box"])
    assert np.array_equal(knn["idx"], g["knn"]["idx"])
    assert np.array_equal(np.where(knn["idx"] >= 0, knn["dist"], 0), np.where(g["knn"]["idx"] >= 0, g["knn"]["dist"], 0))
    m = oracle.match_boxed(dl, dr, qxy, txy, g["box"])
    assert np.array_equal(matches_as_rows(m), g["matches"])
    assert len(m) > 1000
    # the square window is the box {-r, r, -r, r}
    assert np.array_equal(oracle.knn2_boxed(dl, dr, qxy, txy, [-40, 40, -40, 40])["idx"],
                          oracle.knn2(dl, dr, qxy, txy, 40.0)["idx"])


# ---- arbitrary mask (cv knnMatch(mask)) and the depth look-up of the landmark creation ------------------

def test_masked_knn_golden():
    """oracle.knn2_masked against cv2's knnMatch(k=2, mask) (masked_kat.npz): ties, empty and single-entry rows."""
    g = load_golden("masked_kat.npz")
    knn = oracle.knn2_masked(g["q"], g["t"], g["mask"])
    assert np.array_equal(knn["idx"], g["knn"]["idx"])
    assert np.array_equal(np.where(knn["idx"] >= 0, knn["dist"], 0), np.where(g["knn"]["idx"] >= 0, g["knn"]["dist"], 0))
    assert list(knn["idx"][5]) == [-1, -1] and list(knn["idx"][6]) == [123, -1]
    assert np.array_equal(matches_as_rows(oracle.match_from_knn(knn)), g["matches"])
    # the window and box masks are special cases of it
    rng = np.random.default_rng(3)
    qxy = (rng.random((300, 2)) * 200).astype(np.float32)
    txy = (rng.random((700, 2)) * 200).astype(np.float32)
    win = ((np.abs(qxy[:, None, 0] - txy[None, :, 0]) <= 30) & (np.abs(qxy[:, None, 1] - txy[None, :, 1]) <= 30)).astype(np.uint8)
    assert np.array_equal(oracle.knn2_masked(g["q"], g["t"], win), oracle.knn2(g["q"], g["t"], qxy, txy, 30.0))


def test_depth_lookup_restatement():
    """oracle.depth_lookup against an independent numpy statement of tracking.cpp:606-637."""
    rng = np.random.default_rng(8)
    xy = (rng.random((500, 2)) * np.float32([660, 500]) - 10).astype(np.float32)  # some outside the image
    xy[:4] = [[0.49, 0.5], [639.5, 479.49], [639.49, 479.5], [-0.5, 10.0]]
    for dt, img in ((np.uint16, rng.integers(0, 60000, (480, 640)).astype(np.uint16)),
                    (np.float32, (rng.random((480, 640)) * 12).astype(np.float32)),
                    (np.float64, rng.random((480, 640)) * 12)):
        if dt is np.uint16:
            img[rng.random(img.shape) < 0.2] = 0
        got = oracle.depth_lookup(xy, img)
        want = np.zeros(len(xy))
        for i, (x, y) in enumerate(xy.astype(np.float64)):
            u, v = int(x + 0.5), int(y + 0.5)  # C truncation toward zero, like static_cast<int>
            if not (0 <= u < 640 and 0 <= v < 480):
                continue
            m = float(img[v, u]) / 5000.0 if dt is np.uint16 else float(img[v, u])
            if dt is np.uint16 and img[v, u] == 0:
                continue
            if 0.1 <= m <= 10.0:
                want[i] = m
        assert np.array_equal(got, want), dt
        assert (got > 0).sum() > 100
