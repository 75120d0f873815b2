"""Parity of the CUDA path (through the C ABI) against the CPU oracle and the cv2 golden vectors.

Bars (BASELINE.json north_star): FAST keypoint sets, Hamming distances and match indices
bit-exact; Harris/orientation floats within 1e-4 relative (we assert bit equality and fall back to
the stated tolerance only in `test_float_tolerance_statement`); descriptors bit-exact.
Every test here needs a B200 and is marked gpu.
"""
import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest

import oracle
import visionx_slam_b200 as vx
from conftest import assert_kp_equal, load_golden, matches_as_rows
from visionx_slam_b200.synth import sha256, synth_scene, synth_sequence, synth_stereo_sequence

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def h():
    with vx.VxOrb(max_width=1920, max_height=1088, max_batch=8, max_train=20000) as handle:
        yield handle


def _corners(c):
    return np.stack([c["x"], c["y"], c["score"]], 1).astype(np.int32).reshape(-1, 3)


def _knn_equal(a, b):
    return np.array_equal(a["idx"], b["idx"]) and np.array_equal(
        np.where(a["idx"] >= 0, a["dist"], 0), np.where(b["idx"] >= 0, b["dist"], 0))


# ---- stage isolation -----------------------------------------------------------------------------

@pytest.mark.parametrize("shape,dst", [((480, 640), (400, 533)), ((231, 309), (193, 257)), ((376, 1241), (313, 1034)),
                                       ((97, 131), (41, 77)), ((50, 64), (100, 128)), ((33, 7), (33, 5))])
def test_resize_stage(h, shape, dst):
    rng = np.random.default_rng(shape[0] * 7 + dst[1])
    src = rng.integers(0, 256, shape, dtype=np.uint8)
    got = h.debug_resize(src, dst[1], dst[0])
    assert np.array_equal(got, oracle.resize_linear_exact(src, dst[1], dst[0]))


@pytest.mark.parametrize("band_rows", [16, 8, 4])
@pytest.mark.parametrize("shape,dst", [((480, 640), (400, 533)), ((400, 533), (333, 444)), ((278, 370), (231, 309)),
                                       ((161, 214), (134, 179)), ((376, 1241), (313, 1034)), ((1080, 1920), (900, 1600)),
                                       ((97, 131), (49, 77)), ((50, 64), (40, 64)), ((19, 23), (16, 19))])
def test_resize_band_stage(h, shape, dst, band_rows):
    """The production pyramid kernel in isolation, both instantiations: band_rows 16 is resize_band_kernel<true>
    (the batched throughput shape incl. the shared row tails), 8 / 4 resize_band_kernel<false> (latency shape)."""
    rng = np.random.default_rng(shape[0] * 11 + dst[1] + band_rows)
    src = rng.integers(0, 256, shape, dtype=np.uint8)
    got = h.debug_resize_band(src, dst[1], dst[0], band_rows)
    assert np.array_equal(got, oracle.resize_linear_exact(src, dst[1], dst[0]))


BLUR_MARGIN = 13  # kBlurMargin: edge threshold 31 - descriptor window reach 18


def _interior(img, m=BLUR_MARGIN):
    return img[m:img.shape[0] - m, m:img.shape[1] - m]


def test_pyramid_chain_golden(h, golden_c1, monkeypatch):
    """cv2's pyramid levels and blurred levels of the C1 frame.  The production path blurs only what a descriptor
    window can read (BLUR_MARGIN px inside every level); VXORB_FULL_BLUR=1 makes a handle blur whole levels from
    aproned planes — the same kernel — so that cv2's whole-level hashes can be checked, and the interior of the
    production planes must equal the interior of those."""
    g = golden_c1
    a = g["frame_a"]
    h.extract(a)
    monkeypatch.setenv("VXORB_FULL_BLUR", "1")
    with vx.VxOrb() as hf:
        monkeypatch.delenv("VXORB_FULL_BLUR")
        kp, desc = hf.extract(a)
        assert_kp_equal(kp, g["kp_a"])
        assert np.array_equal(desc, g["desc_a"])
        for l in range(8):
            lvl = h.debug_level(0, l, 640, 480)
            assert sha256(lvl) == str(g[f"a_level{l}_sha"]), f"pyramid level {l}"
            full = hf.debug_level(0, l, 640, 480, blurred=True)
            assert sha256(full) == str(g[f"a_blur{l}_sha"]), f"blurred level {l}"
            assert np.array_equal(_interior(h.debug_level(0, l, 640, 480, blurred=True)), _interior(full)), f"level {l}"


@pytest.mark.parametrize("kind,shape,border,thr", [
    ("synth", (480, 640), 3, 20), ("synth", (480, 640), 31, 20), ("noise", (240, 320), 3, 20),
    ("noise", (97, 150), 31, 20), ("noise", (64, 64), 3, 40), ("synth", (134, 179), 3, 7),
    ("noise", (20, 9), 3, 20), ("noise", (7, 7), 3, 20), ("noise", (70, 200), 3, 1)])
def test_fast_stage(h, kind, shape, border, thr):
    rng = np.random.default_rng(shape[0] + shape[1] + border)
    img = synth_scene(shape[1], shape[0], 4) if kind == "synth" else rng.integers(0, 256, shape, dtype=np.uint8)
    got = h.debug_fast(img, thr, border)
    ref = oracle.fast9_nms(img, thr, border)
    assert np.array_equal(_corners(got), _corners(ref))


def test_fast_stage_golden(h, golden_c1):
    g = golden_c1
    levels = oracle.pyramid(g["frame_a"])
    for l in range(8):
        got = h.debug_fast(levels[l], 20, 3)
        assert np.array_equal(_corners(got).astype(np.int16), g[f"a_fast{l}"]), f"FAST level {l}"


@pytest.mark.parametrize("shape", [(480, 640), (134, 179), (65, 64), (31, 200), (5, 9), (1, 40), (40, 1), (3, 3)])
def test_blur_stage(h, shape):
    rng = np.random.default_rng(shape[0] * 3 + shape[1])
    img = rng.integers(0, 256, shape, dtype=np.uint8)
    assert np.array_equal(h.debug_blur(img), oracle.blur7(img))


# ---- whole extraction ----------------------------------------------------------------------------

def test_extract_c1_golden(h, golden_c1):
    g = golden_c1
    for tag in "ab":
        kp, desc = h.extract(g["frame_" + tag])
        assert_kp_equal(kp, g["kp_" + tag])
        assert np.array_equal(desc, g["desc_" + tag])
        kp2, desc2 = h.extract(g["frame_" + tag])  # idempotence / determinism
        assert np.array_equal(kp, kp2) and np.array_equal(desc, desc2)


@pytest.mark.parametrize("name", ["c3_1241x376.npz", "c4_1920x1080.npz"])
def test_extract_large_shapes(name):
    g = load_golden(name)
    w, hh, n = int(g["w"]), int(g["h"]), int(g["n_features"])
    img = synth_scene(w, hh, int(g["seed"]))
    with vx.VxOrb(n_features=n, max_width=1920, max_height=1088) as hx:
        kp, desc = hx.extract(img)
    okp, odesc = oracle.orb_extract(img, n)
    assert_kp_equal(kp, okp)
    assert np.array_equal(desc, odesc)
    if sha256(img) == str(g["frame_sha"]):
        assert_kp_equal(kp, g["kp"])
        assert np.array_equal(desc, g["desc"])


@pytest.mark.parametrize("w,hh,n", [(3840, 2160, 8000), (4095, 3071, 20000)])
def test_extract_4k_vs_oracle(w, hh, n):
    """Maximum sizes: 4K frames with per-level quotas far above the in-shared-memory sort capacity."""
    img = synth_scene(w, hh, 5)
    with vx.VxOrb(n_features=n, max_width=w, max_height=hh, max_batch=1, max_keypoints=n + n // 2) as hx:
        kp, desc = hx.extract(img)
    okp, odesc = oracle.orb_extract(img, n)
    assert len(okp) == n
    assert_kp_equal(kp, okp)
    assert np.array_equal(desc, odesc)


@pytest.mark.parametrize("name", ["grid", "noise", "blank", "tiny60", "small80x70", "lowtex", "odd_357x291"])
def test_extract_adversarial(h, golden_adv, name):
    g = golden_adv
    kp, desc = h.extract(g[name + "_img"])
    assert_kp_equal(kp, g[name + "_kp"])
    assert np.array_equal(desc, g[name + "_desc"])


@pytest.mark.parametrize("tag", ["p500_4", "p300_s15_5", "p2000_8"])
def test_extract_other_params(golden_adv, tag):
    g = golden_adv
    n, sf, nl = g[tag + "_params"]
    img = synth_scene(640, 480, int(g["params_img_seed"]))
    with vx.VxOrb(n_features=int(n), scale_factor=float(sf), n_levels=int(nl)) as hx:
        kp, desc = hx.extract(img)
    assert_kp_equal(kp, g[tag + "_kp"])
    assert np.array_equal(desc, g[tag + "_desc"])


def test_extract_bgr(golden_adv):
    g = golden_adv
    with vx.VxOrb(n_features=500, max_width=320, max_height=240) as hx:
        kp, desc = hx.extract(g["bgr_img"])
    assert_kp_equal(kp, g["bgr_kp"])
    assert np.array_equal(desc, g["bgr_desc"])


def test_extract_strided_input(h):
    big = synth_scene(700, 500, 12)
    view = big[10:490, 30:670]  # 480x640 window with a 700-byte row stride
    kp, desc = h.extract(view)
    okp, odesc = oracle.orb_extract(np.ascontiguousarray(view))
    assert_kp_equal(kp, okp)
    assert np.array_equal(desc, odesc)


def test_extract_batch_equals_single(h):
    frames = [synth_scene(640, 480, s) for s in (31, 32, 33, 34, 35)]
    outs = h.extract_batch(frames)
    for f, (kp, desc) in zip(frames, outs):
        okp, odesc = oracle.orb_extract(f)
        assert_kp_equal(kp, okp)
        assert np.array_equal(desc, odesc)


def test_extract_many_random_frames(h):
    """Wider sample: 12 seeds x 2 shapes, oracle comparison."""
    for seed in range(40, 52):
        for (w, hh) in ((640, 480), (413, 297)):
            img = synth_scene(w, hh, seed)
            kp, desc = h.extract(img)
            okp, odesc = oracle.orb_extract(img)
            assert_kp_equal(kp, okp)
            assert np.array_equal(desc, odesc), f"seed {seed} {w}x{hh}"


def test_extract_random_shapes_and_params():
    """Randomised frame sizes (40..900 x 40..700), feature counts, level counts and scale factors, scene / noise /
    gradient content, batched and streamed: every output equals the oracle's (scratch/soak.py is the long form:
    600 cases, 0 mismatches)."""
    rng = np.random.default_rng(2026)
    for case in range(14):
        w, hh = int(rng.integers(40, 900)), int(rng.integers(40, 700))
        nf, nl = int(rng.choice([50, 300, 1000, 2500])), int(rng.integers(1, 10))
        sf = float(rng.choice([1.1, 1.2, 1.3, 1.5, 2.0]))
        kind = rng.choice(["scene", "noise", "grad"])
        if kind == "scene":
            img = synth_scene(w, hh, int(rng.integers(0, 1000)))
        elif kind == "noise":
            img = rng.integers(0, 256, (hh, w), dtype=np.uint8)
        else:
            img = ((np.add.outer(np.arange(hh), np.arange(w)) * 3) % 256).astype(np.uint8)
        frames = [img, np.roll(img, 3, axis=1)]
        with vx.VxOrb(n_features=nf, scale_factor=sf, n_levels=nl, max_width=w, max_height=hh, max_batch=2,
                      max_keypoints=nf * 3 + 512) as hx:
            outs = hx.extract_batch(frames)
            tracked = hx.track(np.stack(frames), 24.0)
        for f, (kp, desc), (tkp, tdesc, _) in zip(frames, outs, tracked):
            okp, odesc = oracle.orb_extract(f, nf, sf, nl)
            assert_kp_equal(kp, okp)
            assert np.array_equal(desc, odesc), f"case {case}: {w}x{hh} n={nf} levels={nl} sf={sf} {kind}"
            assert_kp_equal(tkp, okp)
            assert np.array_equal(tdesc, odesc)


def test_extract_capacity_error(golden_c1):
    with vx.VxOrb(max_keypoints=300) as hx:
        kp, desc, rc, n = hx.extract(golden_c1["frame_a"], return_code=True)
    assert rc == vx.E_CAPACITY and n == 1000 and len(kp) == 300
    assert np.array_equal(kp, golden_c1["kp_a"][:300])  # truncated prefix is still canonical order
    assert np.array_equal(desc, golden_c1["desc_a"][:300])


def test_float_tolerance_statement(h, golden_c1):
    """north_star tolerance: Harris / orientation within 1e-4 relative (we are bit-equal)."""
    kp, _ = h.extract(golden_c1["frame_a"])
    ref = golden_c1["kp_a"]
    np.testing.assert_allclose(kp["response"], ref["response"], rtol=1e-4)
    np.testing.assert_allclose(kp["angle"], ref["angle"], rtol=1e-4)


# ---- matching ------------------------------------------------------------------------------------

def test_match_c1_golden(h, golden_c1):
    g = golden_c1
    da, db = g["desc_a"], g["desc_b"]
    assert _knn_equal(h.knn2(da, db), g["knn"])
    assert np.array_equal(matches_as_rows(h.match(da, db)), g["matches"])
    qxy = np.stack([g["kp_a"]["x"], g["kp_a"]["y"]], 1)
    txy = np.stack([g["kp_b"]["x"], g["kp_b"]["y"]], 1)
    assert _knn_equal(h.knn2(da, db, qxy, txy, 32.0), g["knn_win32"])
    assert np.array_equal(matches_as_rows(h.match(da, db, 0.8, qxy, txy, 32.0)), g["matches_win32"])
    assert np.array_equal(matches_as_rows(h.match_crosscheck(da, db)), g["crosscheck"])


@pytest.mark.parametrize("tag", ["ties", "nt1", "nt2", "rand", "dup"])
def test_match_kat(h, golden_kat, tag):
    g = golden_kat
    q, t = g[tag + "_q"], g[tag + "_t"]
    assert _knn_equal(h.knn2(q, t), g[tag + "_knn"])
    assert np.array_equal(matches_as_rows(h.match(q, t)), g[tag + "_matches"])
    assert np.array_equal(matches_as_rows(h.match_crosscheck(q, t)), g[tag + "_crosscheck"])


def test_match_windowed_kat(h, golden_kat):
    g = golden_kat
    knn = h.knn2(g["dup_q"], g["dup_t"], g["dup_qxy"], g["dup_txy"], 64.0)
    assert _knn_equal(knn, g["dup_knn_win64"])
    m = h.match(g["dup_q"], g["dup_t"], 0.8, g["dup_qxy"], g["dup_txy"], 64.0)
    assert np.array_equal(matches_as_rows(m), g["dup_matches_win64"])


@pytest.mark.parametrize("nq,nt,radius,kind", [
    (1000, 1000, 32.0, "image"), (1000, 1000, 16.0, "image"), (997, 1003, 100.0, "image"),
    (300, 2500, 32.0, "cluster"), (2000, 2000, 24.0, "lattice"), (500, 700, 32.0, "far"),
    (64, 1, 32.0, "image"), (40, 6000, 32.0, "image"),
])
def test_match_windowed_vs_oracle(h, nq, nt, radius, kind):
    """Cell-binned windowed 2-NN (few distinct descriptors -> heavy distance ties) against the
    oracle's masked knn: positions on a lattice put many pairs exactly on the window edge."""
    rng = np.random.default_rng(nq * 7 + nt)
    base = rng.integers(0, 256, (24, 32), dtype=np.uint8)
    t = base[rng.integers(0, 24, nt)].copy()
    q = base[rng.integers(0, 24, nq)].copy()
    t[::3, rng.integers(0, 32)] ^= 0x5a
    q[::5, rng.integers(0, 32)] ^= 0x0f
    if kind == "image":
        txy = rng.random((nt, 2), np.float32) * np.float32([640, 480])
        qxy = rng.random((nq, 2), np.float32) * np.float32([640, 480])
    elif kind == "cluster":
        txy = (rng.normal(0, 20, (nt, 2)) + [300, 200]).astype(np.float32)
        qxy = (rng.normal(0, 40, (nq, 2)) + [300, 200]).astype(np.float32)
    elif kind == "lattice":
        txy = (rng.integers(0, 40, (nt, 2)) * 8).astype(np.float32)      # |dx| == radius happens often
        qxy = (rng.integers(0, 40, (nq, 2)) * 8).astype(np.float32)
    else:  # negative and large coordinates, queries partly outside the train bounding box
        txy = (rng.random((nt, 2), np.float32) * 2000 - 1000).astype(np.float32) + np.float32(1.0e5)
        qxy = (rng.random((nq, 2), np.float32) * 2400 - 1200).astype(np.float32) + np.float32(1.0e5)
    assert _knn_equal(h.knn2(q, t, qxy, txy, radius), oracle.knn2(q, t, qxy, txy, radius))
    assert np.array_equal(matches_as_rows(h.match(q, t, 0.8, qxy, txy, radius)),
                          matches_as_rows(oracle.match(q, t, 0.8, qxy, txy, radius)))


def test_match_masked_golden(h):
    """Arbitrary byte mask = cv knnMatch(k=2, mask) (masked_kat.npz: ties, empty and single-entry rows), then larger
    random masks against the oracle, and the window mask written out as bytes against the windowed kernel."""
    g = load_golden("masked_kat.npz")
    assert _knn_equal(h.knn2_masked(g["q"], g["t"], g["mask"]), g["knn"])
    assert np.array_equal(matches_as_rows(h.match_masked(g["q"], g["t"], g["mask"])), g["matches"])
    rng = np.random.default_rng(31)
    for nq, nt, density in ((1000, 1000, 0.05), (257, 3000, 0.5), (3, 5, 1.0), (700, 129, 0.002)):
        t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
        q = t[rng.integers(0, nt, nq)].copy()
        q[:, 7] ^= rng.integers(0, 8, nq).astype(np.uint8)
        mask = (rng.random((nq, nt)) < density).astype(np.uint8) * 3
        assert _knn_equal(h.knn2_masked(q, t, mask), oracle.knn2_masked(q, t, mask)), (nq, nt)
        assert np.array_equal(matches_as_rows(h.match_masked(q, t, mask, 0.7)),
                              matches_as_rows(oracle.match_from_knn(oracle.knn2_masked(q, t, mask), 0.7)))
    qxy = (rng.random((700, 2)) * 300).astype(np.float32)
    txy = (rng.random((129, 2)) * 300).astype(np.float32)
    win = ((np.abs(qxy[:, None, 0] - txy[None, :, 0]) <= 40) & (np.abs(qxy[:, None, 1] - txy[None, :, 1]) <= 40)).astype(np.uint8)
    assert _knn_equal(h.knn2_masked(q, t, win), h.knn2(q, t, qxy, txy, 40.0))


def test_depth_lookup_vs_oracle(golden_c1):
    """Tracking::CreateLandmarksFromDepth's per-feature depth (tracking.cpp:606-637) for a resident frame, all three
    depth image types, against the oracle."""
    rng = np.random.default_rng(5)
    with vx.VxOrb() as hx:
        kp, desc, fid = hx.extract_id(golden_c1["frame_a"])
        xy = np.stack([kp["x"], kp["y"]], 1)
        d16 = rng.integers(0, 60000, (480, 640)).astype(np.uint16)
        d16[rng.random(d16.shape) < 0.2] = 0
        for depth in (d16, (rng.random((480, 640)) * 12).astype(np.float32), rng.random((480, 640)) * 12):
            got = hx.depth_lookup(fid, depth)
            assert np.array_equal(got, oracle.depth_lookup(xy, depth)), depth.dtype
            assert (got > 0).sum() > 300
        small = np.full((100, 100), 5000, np.uint16)  # a depth image smaller than the frame: features outside are skipped
        got = hx.depth_lookup(fid, small)
        assert np.array_equal(got, oracle.depth_lookup(xy, small)) and 0 < (got > 0).sum() < len(kp)
        with pytest.raises(vx.VxOrbError):
            hx.depth_lookup(fid + 100, d16)


def test_match_empty_sides(h):
    z = np.zeros((0, 32), np.uint8)
    q = np.zeros((5, 32), np.uint8)
    assert len(h.match(z, q)) == 0 and len(h.match(q, z)) == 0  # orb_matcher.cpp:18-20


@pytest.mark.parametrize("nq,nt", [(1, 1), (1, 300), (129, 257), (1000, 1000), (2000, 2000), (4000, 16384)])
def test_match_sizes_vs_oracle(h, nq, nt):
    rng = np.random.default_rng(nq + nt)
    t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    k = min(nq, nt)
    q[:k:2] = t[rng.permutation(nt)[:len(q[:k:2])]]  # plant exact and near duplicates
    q[:k:4, 5] ^= 0x11
    assert _knn_equal(h.knn2(q, t), oracle.knn2(q, t))
    for ratio in (0.8, 0.6, 0.95):
        assert np.array_equal(matches_as_rows(h.match(q, t, ratio)), matches_as_rows(oracle.match(q, t, ratio)))
    assert np.array_equal(matches_as_rows(h.match_crosscheck(q, t)), matches_as_rows(oracle.match_crosscheck(q, t)))
    m = oracle.match(q, t, 0.8)
    assert np.array_equal(matches_as_rows(h.match_filtered(q, t)), matches_as_rows(oracle.postfilter(m)))


def test_match_properties_full_size(h):
    """Size-independent properties at C4 size: self-match, distance = popcount, ordering."""
    rng = np.random.default_rng(99)
    t = rng.integers(0, 256, (16384, 32), dtype=np.uint8)
    q = t[:4000]
    knn = h.knn2(q, t)
    assert np.array_equal(knn["idx"][:, 0], np.arange(4000)) and not knn["dist"][:, 0].any()
    d1 = np.unpackbits(q ^ t[knn["idx"][:, 1]], axis=1).sum(1)
    assert np.array_equal(d1, knn["dist"][:, 1])
    m = h.match(q, t, 0.8)
    assert np.all(np.diff(m["query_idx"]) > 0)


# ---- streaming sequence path (BASELINE config 2) -------------------------------------------------

def _oracle_sequence(frames, radius):
    out = []
    prev = None
    for f in frames:
        kp, desc = oracle.orb_extract(f)
        if prev is None:
            m = np.zeros(0, oracle.MATCH_DTYPE)
        else:
            pk, pd = prev
            m = oracle.match(pd, desc, 0.8, np.stack([pk["x"], pk["y"]], 1) if radius > 0 else None,
                             np.stack([kp["x"], kp["y"]], 1) if radius > 0 else None, radius)
        out.append((kp, desc, m))
        prev = (kp, desc)
    return out


@pytest.mark.parametrize("radius", [32.0, 0.0])
def test_track_sequence_chunks(radius):
    frames = synth_sequence(640, 480, 11, seed=3)
    ref = _oracle_sequence(frames, radius)
    with vx.VxOrb(max_batch=4) as hx:
        got = []
        for i in range(0, 11, 4):  # chunks of 4, 4, 3: carry crosses two chunk boundaries and 3 slots
            got += hx.track(frames[i:i + 4], radius)
        for i, ((kp, desc, m), (okp, odesc, om)) in enumerate(zip(got, ref)):
            assert_kp_equal(kp, okp)
            assert np.array_equal(desc, odesc)
            assert np.array_equal(matches_as_rows(m), matches_as_rows(om)), f"frame {i}"
        assert len(got[1][2]) > 200  # the sequence really matches frame to frame
        hx.track_reset()
        again = hx.track(frames[4:6], radius)
        assert len(again[0][2]) == 0  # first frame after a reset has no predecessor
        assert np.array_equal(matches_as_rows(again[1][2]), matches_as_rows(ref[5][2]))


def test_track_tie_tails():
    """Response ties push the keypoint count above n_features (cv::ORB keeps every tie): the streaming path copies
    only the first n_features rows of every frame with the chunk and fetches such tails in vxorb_track_wait —
    keypoints, descriptors, matches and matched points beyond row n_features must still arrive."""
    rng = np.random.default_rng(5)
    patch = rng.integers(0, 256, (14, 14)).astype(np.uint8)
    img = np.full((240, 320), 128, np.uint8)
    for y in range(36, 190, 24):
        for x in range(36, 270, 24):
            img[y:y + 14, x:x + 14] = patch        # identical copies: identical scores and responses
    frames = np.stack([img, np.roll(img, 2, axis=1), np.roll(img, 4, axis=1)])
    for n_features, n_levels in ((50, 1), (100, 2)):
        ext = [oracle.orb_extract(f, n_features, 1.2, n_levels) for f in frames]
        assert min(len(k) for k, _ in ext) > n_features  # the tails exist
        with vx.VxOrb(n_features=n_features, n_levels=n_levels, max_width=320, max_height=240, max_batch=3) as hx:
            assert hx.kp_cap >= max(len(k) for k, _ in ext)
            for radius in (0.0, 16.0):
                hx.track_reset()
                got = hx.track_info(frames, radius)
                prev = None
                for i, ((kp, desc, m, info, pts), (okp, odesc)) in enumerate(zip(got, ext)):
                    assert_kp_equal(kp, okp)
                    assert np.array_equal(desc, odesc)
                    if prev is not None:
                        pk, pd = prev
                        qxy, txy = np.stack([pk["x"], pk["y"]], 1), np.stack([okp["x"], okp["y"]], 1)
                        om = oracle.match(pd, odesc, 0.8, qxy if radius > 0 else None, txy if radius > 0 else None, radius)
                        assert np.array_equal(matches_as_rows(m), matches_as_rows(om)), f"frame {i} radius {radius}"
                        opts, _ = oracle.correspondences(qxy, txy, om)
                        assert np.array_equal(pts, opts)
                    prev = (okp, odesc)


def test_track_bgr_chunks():
    """Colour chunks through the streaming path (what System::Run feeds, system.cpp:46): converted on the device,
    then identical to streaming the oracle's gray conversion; a strided (padded-row) chunk gives the same."""
    rng = np.random.default_rng(12)
    gray = synth_sequence(320, 240, 5, seed=8)
    # a real colour sequence: three differently weighted / offset copies of the scene, not gray replicated
    bgr = np.stack([gray, np.roll(gray, 2, axis=2), 255 - gray], axis=-1).astype(np.uint8)
    bgr[..., 1] = np.clip(bgr[..., 1].astype(np.int32) + rng.integers(-6, 7, gray.shape), 0, 255).astype(np.uint8)
    conv = np.stack([oracle.bgr2gray(f) for f in bgr])
    with vx.VxOrb(n_features=500, max_width=320, max_height=240, max_batch=5) as hx:
        want = hx.track(conv, 24.0)
        hx.track_reset()
        got = hx.track_bgr(bgr, 24.0)
        hx.track_reset()
        padded = np.zeros((5, 240, 336, 3), np.uint8)
        padded[:, :, :320] = bgr
        got_strided = hx.track_bgr(padded[:, :, :320], 24.0)  # rows 1008 B apart: the strided upload path
    okp, odesc = oracle.orb_extract(conv[1], 500)
    assert_kp_equal(want[1][0], okp)
    for (kp, desc, m), (wkp, wdesc, wm) in zip(got, want):
        assert_kp_equal(kp, wkp)
        assert np.array_equal(desc, wdesc)
        assert np.array_equal(matches_as_rows(m), matches_as_rows(wm))
    for (kp, desc, m), (wkp, wdesc, wm) in zip(got_strided, want):
        assert_kp_equal(kp, wkp)
        assert np.array_equal(desc, wdesc)
        assert np.array_equal(matches_as_rows(m), matches_as_rows(wm))
    assert len(got[1][2]) > 100


def test_track_pipelined_slots():
    """Three chunks in flight on the three pipeline slots, pinned buffers, waits in order."""
    frames = synth_sequence(640, 480, 9, seed=5)
    ref = _oracle_sequence(frames, 32.0)
    with vx.VxOrb(max_batch=3) as hx:
        cap = hx.kp_cap
        pin_in = vx.PinnedArray((9, 480, 640), np.uint8)
        pin_in.array[:] = frames
        outs = []
        slots = []
        for c in range(3):
            o = dict(kp=vx.PinnedArray((3, cap), vx.KEYPOINT_DTYPE), desc=vx.PinnedArray((3, cap, 32), np.uint8),
                     nk=vx.PinnedArray((3,), np.int32), mt=vx.PinnedArray((3, cap), vx.MATCH_DTYPE),
                     nm=vx.PinnedArray((3,), np.int32))
            outs.append(o)
            slots.append(hx.track_submit(pin_in.array[3 * c:3 * c + 3], 32.0, o["kp"].array, o["desc"].array,
                                         o["nk"].array, o["mt"].array, o["nm"].array))
        assert sorted(slots) == [0, 1, 2]
        for s in slots:
            hx.track_wait(s)
        for c in range(3):
            o = outs[c]
            for j in range(3):
                okp, odesc, om = ref[3 * c + j]
                n, nm = int(o["nk"].array[j]), int(o["nm"].array[j])
                assert_kp_equal(o["kp"].array[j, :n], okp)
                assert np.array_equal(o["desc"].array[j, :n], odesc)
                assert np.array_equal(matches_as_rows(o["mt"].array[j, :nm]), matches_as_rows(om))


def test_resident_equals_track():
    frames = synth_sequence(640, 480, 6, seed=8)
    with vx.VxOrb(max_batch=6) as hx:
        t = hx.track(frames, 32.0)
        hx.resident_upload(frames)
        hx.resident_run(32.0)
        r = hx.resident_fetch()
        for (a, b, c), (d, e, f) in zip(t, r):
            assert np.array_equal(a, d) and np.array_equal(b, e) and np.array_equal(c, f)
        hx.resident_run(32.0)  # second run: frame 0 now matches against the carried last frame
        r2 = hx.resident_fetch()
        assert len(r2[0][2]) >= 0 and np.array_equal(r2[1][2], r[1][2])
        st = hx.stage_times()
        assert hx.launch_count() > 0 and st["fast"][1] > 0


def test_resident_passes_pipelined_and_repeatable():
    """Seven resident passes back to back (they rotate over the pipeline slots and overlap on the device, the
    blur on side streams): every pass from the second on sees the same carry, so its output must equal the
    oracle's for [last frame, frame 0, ...] and be bit-identical from pass to pass; the instrumented serial
    pass must give the same bytes as the overlapped one."""
    frames = synth_sequence(640, 480, 5, seed=12)
    ref = _oracle_sequence(np.concatenate([frames[-1:], frames]), 32.0)[1:]  # frame 0 matched against the carried last frame
    with vx.VxOrb(max_batch=5) as hx:
        hx.resident_upload(frames)
        hx.resident_run(32.0)
        fetched = []
        for _ in range(6):
            hx.resident_run(32.0)
        fetched.append(hx.resident_fetch())
        hx.resident_run(32.0)
        fetched.append(hx.resident_fetch())
        hx.stage_times_reset(True)   # serial, instrumented schedule
        hx.resident_run(32.0)
        fetched.append(hx.resident_fetch())
        hx.stage_times_reset(False)
        for r in fetched:
            for (kp, desc, m), (okp, odesc, om) in zip(r, ref):
                assert_kp_equal(kp, okp)
                assert np.array_equal(desc, odesc)
                assert np.array_equal(matches_as_rows(m), matches_as_rows(om))


# ---- the batch shapes the benchmark times (batch >= 8 switches kernel instantiations) ------------------

def _oracle_extract_many(frames, n_features=1000):
    with ThreadPoolExecutor(max(1, len(os.sched_getaffinity(0)))) as ex:  # the C oracle releases the GIL
        return list(ex.map(lambda f: oracle.orb_extract(f, n_features), frames))


def _oracle_pairs(ext, radius, first_prev=None):
    """Frame-to-frame matches of consecutive oracle extractions; first_prev = the carried frame before frame 0."""
    out, prev = [], first_prev
    for kp, desc in ext:
        if prev is None:
            out.append(np.zeros(0, oracle.MATCH_DTYPE))
        else:
            pk, pd = prev
            win = radius > 0
            out.append(oracle.match(pd, desc, 0.8, np.stack([pk["x"], pk["y"]], 1) if win else None,
                                    np.stack([kp["x"], kp["y"]], 1) if win else None, radius))
        prev = (kp, desc)
    return out


@pytest.mark.parametrize("batch", [16, 128])
def test_bench_batch_shapes_vs_oracle(batch):
    """Batches of 16 and 128 frames of 640x480 — the shapes bench.py times: resize_band_kernel<true>, describe with 16
    keypoints per warp, FAST / blur / select over a batched grid — through extract_batch, track (window and brute
    force) and the resident path, every frame's keypoints, descriptors and matches against the oracle bit for bit;
    all 8 pyramid and blurred levels of the first, a middle and the last frame against the oracle's."""
    frames = synth_sequence(640, 480, batch, seed=21)
    ext = _oracle_extract_many(frames)
    with vx.VxOrb(max_batch=batch) as hx:
        outs = hx.extract_batch(frames)
        for i, ((kp, desc), (okp, odesc)) in enumerate(zip(outs, ext)):
            assert_kp_equal(kp, okp)
            assert np.array_equal(desc, odesc), f"extract_batch frame {i}"
        for f in (0, batch // 2 - 1, batch - 1):
            levels = oracle.pyramid(frames[f])
            for l in range(8):
                assert np.array_equal(hx.debug_level(f, l, 640, 480), levels[l]), f"frame {f} level {l}"
                assert np.array_equal(_interior(hx.debug_level(f, l, 640, 480, blurred=True)),
                                      _interior(oracle.blur7(levels[l]))), f"frame {f} blurred level {l}"
        for radius in (32.0, 0.0):
            ref_m = _oracle_pairs(ext, radius)
            hx.track_reset()
            got = hx.track(frames, radius)
            for i, ((kp, desc, m), (okp, odesc)) in enumerate(zip(got, ext)):
                assert_kp_equal(kp, okp)
                assert np.array_equal(desc, odesc)
                assert np.array_equal(matches_as_rows(m), matches_as_rows(ref_m[i])), f"track r={radius} frame {i}"
            # resident passes: from the second pass on, frame 0 is matched against the carried last frame
            ref_c = _oracle_pairs(ext, radius, first_prev=ext[-1])
            hx.resident_upload(frames)
            for _ in range(4):  # rotates over all pipeline slots
                hx.resident_run(radius)
            res = hx.resident_fetch()
            for i, ((kp, desc, m), (okp, odesc)) in enumerate(zip(res, ext)):
                assert_kp_equal(kp, okp)
                assert np.array_equal(desc, odesc)
                assert np.array_equal(matches_as_rows(m), matches_as_rows(ref_c[i])), f"resident r={radius} frame {i}"


def test_pipelined_passes_are_deterministic():
    """Race check by repetition (compute-sanitizer racecheck is closed on this GPU pool): 60 resident passes of a
    16-frame chunk, pipelined over the slots with the blur on side streams, windowed and brute-force matchers in
    turn; every pass of the same matcher must produce the same bytes (FAST's double-buffered tiles and score maps,
    describe's recycled cp.async windows and the carry between slots are the places a race would show)."""
    import hashlib
    frames = synth_sequence(640, 480, 16, seed=77)
    with vx.VxOrb(max_batch=16) as hx:
        hx.resident_upload(frames)
        for radius in (32.0, 0.0):
            hx.resident_run(radius)
            digests = set()
            for i in range(30):
                hx.resident_run(radius)
                if i % 3 == 2:  # fetch synchronises only every third pass, so passes really overlap
                    res = hx.resident_fetch()
                    digests.add(hashlib.sha256(b"".join(k.tobytes() + d.tobytes() + m.tobytes() for k, d, m in res)).hexdigest())
            assert len(digests) == 1, f"radius {radius}: {len(digests)} different results over 10 fetches"
        okp, odesc = oracle.orb_extract(frames[5])
        assert_kp_equal(res[5][0], okp)
        assert np.array_equal(res[5][1], odesc)


def test_bench_batch_kitti_shape_vs_oracle():
    """1241x376, 2000 keypoints, batch 8 (>= 8: the throughput instantiations on BASELINE config 3's geometry)."""
    frames = synth_sequence(1241, 376, 8, seed=5)
    ext = _oracle_extract_many(frames, 2000)
    ref_m = _oracle_pairs(ext, 48.0)
    with vx.VxOrb(n_features=2000, max_width=1241, max_height=376, max_batch=8, max_train=4096) as hx:
        outs = hx.extract_batch(frames)
        got = hx.track(frames, 48.0)
        for i, ((kp, desc), (tkp, tdesc, m), (okp, odesc)) in enumerate(zip(outs, got, ext)):
            assert_kp_equal(kp, okp)
            assert np.array_equal(desc, odesc), f"frame {i}"
            assert_kp_equal(tkp, okp)
            assert np.array_equal(tdesc, odesc)
            assert np.array_equal(matches_as_rows(m), matches_as_rows(ref_m[i])), f"frame {i}"
        levels = oracle.pyramid(frames[7])
        for l in range(8):
            assert np.array_equal(hx.debug_level(7, l, 1241, 376), levels[l]), f"level {l}"


# ---- the reference's interface, as the reference's caller uses it --------------------------------

def test_reference_interface_flow(golden_c1):
    """tracking.cpp:42,208 — Extract(frame) on two frames, then Match(last, curr, matches)."""
    g = golden_c1
    extractor = vx.ORBExtractor()      # system.cpp:15
    matcher = vx.ORBMatcher()          # system.cpp:16
    last, curr = vx.Frame(0, 0.0, g["frame_a"]), vx.Frame(1, 0.1, g["frame_b"])
    extractor.Extract(last)
    extractor.Extract(curr)
    assert len(last.Features()) == 1000 and last.Descriptors().shape == (1000, 32)
    f0 = last.Features()[0]
    assert f0.position == (float(g["kp_a"]["x"][0]), float(g["kp_a"]["y"][0]))
    assert f0.response == float(g["kp_a"]["response"][0])
    matches = [None]
    n = matcher.Match(last, curr, matches)
    assert n == len(g["matches"]) == len(matches)
    rows = np.array([(m.queryIdx, m.trainIdx, int(m.distance)) for m in matches], np.int32)
    assert np.array_equal(rows, g["matches"])
    blank = vx.Frame(2, 0.2, np.zeros((480, 640), np.uint8))
    extractor.Extract(blank)
    assert len(blank.Features()) == 0 and matcher.Match(last, blank, matches) == 0 and matches == []


def test_reference_interface_flow_resident(golden_c1):
    """The same flow with the matcher sharing the extractor's handle: Match runs on the descriptors Extract left
    on the device (no re-upload); a frame that fell out of the 8-entry ring silently takes the upload path."""
    g = golden_c1
    extractor = vx.ORBExtractor()
    matcher = vx.ORBMatcher(extractor=extractor)
    last, curr = vx.Frame(0, 0.0, g["frame_a"]), vx.Frame(1, 0.1, g["frame_b"])
    extractor.Extract(last)
    extractor.Extract(curr)
    matches = []
    assert matcher.Match(last, curr, matches) == len(g["matches"])
    rows = np.array([(m.queryIdx, m.trainIdx, int(m.distance)) for m in matches], np.int32)
    assert np.array_equal(rows, g["matches"]) and (matcher.resident_calls, matcher.upload_calls) == (1, 0)
    for i in range(8):  # push `last` out of the ring
        extractor.Extract(vx.Frame(10 + i, 1.0, g["frame_b"]))
    assert extractor.resident_id(last) == 0
    assert matcher.Match(last, curr, matches) == len(g["matches"]) and matcher.upload_calls == 1
    rows = np.array([(m.queryIdx, m.trainIdx, int(m.distance)) for m in matches], np.int32)
    assert np.array_equal(rows, g["matches"])


def test_reference_interface_grows_to_the_frame_size(golden_c1):
    """cv::ORB takes frames of any size: an extractor built for 96 x 96 rebuilds its handle on the first larger
    frame; ids of the old handle are void (the pair straddling the rebuild takes the upload path), later pairs are
    matched from the device again, and every result equals the golden one."""
    g = golden_c1
    small = np.ascontiguousarray(g["frame_a"][:90, :90])
    extractor = vx.ORBExtractor(max_width=96, max_height=96)
    matcher = vx.ORBMatcher(extractor=extractor)
    tiny = vx.Frame(0, 0.0, small)
    extractor.Extract(tiny)
    assert extractor.rebuilds == 0
    okp, odesc = oracle.orb_extract(small)
    assert len(tiny.Features()) == len(okp) and np.array_equal(tiny.Descriptors(), odesc)
    last, curr = vx.Frame(1, 0.1, g["frame_a"]), vx.Frame(2, 0.2, g["frame_b"])
    extractor.Extract(last)
    assert extractor.rebuilds == 1 and extractor.resident_id(tiny) == 0 and extractor.resident_id(last) != 0
    extractor.Extract(curr)
    assert extractor.rebuilds == 1
    assert_kp_equal(last._keypoints, g["kp_a"])
    assert np.array_equal(last.Descriptors(), g["desc_a"]) and np.array_equal(curr.Descriptors(), g["desc_b"])
    matches = []
    assert matcher.Match(last, curr, matches) == len(g["matches"]) and matcher.resident_calls == 1
    rows = np.array([(m.queryIdx, m.trainIdx, int(m.distance)) for m in matches], np.int32)
    assert np.array_equal(rows, g["matches"])
    if len(odesc):  # the frame of the old handle is still matchable: its descriptors go up from the host
        n = matcher.Match(tiny, curr, matches)
        assert n == len(oracle.match(odesc, g["desc_b"], 0.8)) and matcher.upload_calls == 1


def test_match_ids_vs_oracle(golden_c1):
    """vxorb_extract_id + vxorb_match_ids (ratio, post-filter, cross-check, window) and an uploaded map set."""
    g = golden_c1
    with vx.VxOrb(max_train=64) as hx:  # max_train is only a sizing hint: the scratch grows
        ka, da, ia = hx.extract_id(g["frame_a"])
        kb, db, ib, info = hx.extract_id(g["frame_b"], with_info=True)
        assert_kp_equal(ka, g["kp_a"])
        assert_kp_equal(kb, g["kp_b"])
        assert np.array_equal(da, g["desc_a"]) and np.array_equal(db, g["desc_b"])
        assert ib == ia + 1 and hx.set_size(ia) == len(da) and info["n_keypoints"] == len(kb)
        assert np.array_equal(matches_as_rows(hx.match_ids(ia, ib)), g["matches"])
        assert np.array_equal(matches_as_rows(hx.match_ids(ia, ib, mode=2)), g["crosscheck"])
        assert np.array_equal(matches_as_rows(hx.match_ids(ia, ib, radius=32.0)), g["matches_win32"])
        om = oracle.match(da, db, 0.8)
        assert np.array_equal(matches_as_rows(hx.match_ids(ia, ib, mode=1)), matches_as_rows(oracle.postfilter(om)))
        assert np.array_equal(matches_as_rows(hx.match_ids(ib, ia, nn_ratio=0.6)), matches_as_rows(oracle.match(db, da, 0.6)))
        # a 16 384-row map kept on the device (BASELINE config 4's shape), matched against a frame id
        rng = np.random.default_rng(4)
        world = rng.integers(0, 256, (16384, 32), dtype=np.uint8)
        world[::7][:len(da)] = da[:len(world[::7])]
        sid = hx.set_upload(world)
        assert hx.set_size(sid) == 16384
        assert np.array_equal(matches_as_rows(hx.match_ids(ia, sid)), matches_as_rows(oracle.match(da, world, 0.8)))
        with pytest.raises(vx.VxOrbError):
            hx.match_ids(ia, sid, radius=16.0)  # the map has no positions
        hx.set_free(sid)
        with pytest.raises(vx.VxOrbError):
            hx.match_ids(ia, sid)
        for _ in range(8):
            hx.extract_id(g["frame_b"])
        with pytest.raises(vx.VxOrbError):  # overwritten: the caller falls back to the host-buffer call
            hx.match_ids(ia, ib)
        # a different geometry re-captures the per-entry graphs; ids of the old geometry stay valid (descriptors only)
        small = synth_scene(320, 240, 3)
        ks, ds, is_ = hx.extract_id(small)
        okp, odesc = oracle.orb_extract(small)
        assert_kp_equal(ks, okp)
        assert np.array_equal(ds, odesc)
        k2, d2, i2 = hx.extract_id(g["frame_b"])
        assert np.array_equal(matches_as_rows(hx.match_ids(is_, i2)), matches_as_rows(oracle.match(odesc, g["desc_b"], 0.8)))


# ---- caller-side per-frame quantities ----------------------------------------

def _check_info(info, gray, kp, w, h):
    mean, sd = oracle.image_stats(gray)
    assert info["mean"] == mean and info["stddev"] == sd  # integer sums + OpenCV's double expression: bit-exact
    occ, mask = oracle.grid_coverage(np.stack([kp["x"], kp["y"]], 1) if len(kp) else None, w, h)
    assert info["grid_occupied"] == occ and info["grid_mask"] == mask
    assert info["n_keypoints"] == len(kp)


@pytest.mark.parametrize("name", ["synth", "noise_odd", "const", "bright", "dark_w3", "bgr"])
def test_extract_info_golden(h, name):
    """Image statistics against cv2.meanStdDev (golden), grid coverage against the oracle, single frames."""
    g = load_golden("caller_kat.npz")
    img = g[name + "_img"]
    kp, desc, info = h.extract_info(img)
    assert info["mean"] == g[name + "_stats"][0] and info["stddev"] == g[name + "_stats"][1]
    gray = oracle.bgr2gray(img) if img.ndim == 3 else img
    okp, odesc = oracle.orb_extract(gray)
    assert_kp_equal(kp, okp)
    assert np.array_equal(desc, odesc)
    _check_info(info, gray, kp, img.shape[1], img.shape[0])
    assert info["n_matches"] == 0 and info["parallax"] == 0.0


@pytest.mark.parametrize("size,radius", [((640, 480), 32.0), ((357, 291), 0.0)])
def test_track_info_vs_oracle(size, radius):
    """Streaming path with the per-frame info and matched positions, chunks of 3 + 2 frames."""
    w, hh = size
    frames = synth_sequence(w, hh, 5, seed=8)
    ref = _oracle_sequence(frames, radius)
    with vx.VxOrb(max_width=w, max_height=hh, max_batch=3) as hx:
        got = hx.track_info(frames[:3], radius) + hx.track_info(frames[3:], radius)
    prev_xy = None
    for i, ((kp, desc, m, info, pts), (okp, odesc, om)) in enumerate(zip(got, ref)):
        assert_kp_equal(kp, okp)
        assert np.array_equal(matches_as_rows(m), matches_as_rows(om)), f"frame {i}"
        _check_info(info, frames[i], kp, w, hh)
        xy = np.stack([kp["x"], kp["y"]], 1)
        assert info["n_matches"] == len(om)
        if i == 0:
            assert info["parallax"] == 0.0 and len(pts) == 0
        else:
            opts, opar = oracle.correspondences(prev_xy, xy, om)
            assert np.array_equal(pts, opts)
            # distances are computed un-fused in double on both sides; the GPU sums them with a fixed tree,
            # the reference serially: 1e-12 relative covers the re-association
            assert abs(info["parallax"] - opar) <= 1e-12 * max(1.0, abs(opar))
            assert opar > 0.5  # the sequence really moves
        prev_xy = xy


# ---- BASELINE config 3: stereo pair, left -> right matching along the epipolar band ------------------

def test_stereo_pair_boxed_golden():
    """1241x376 rectified pair, 2000 kp per view, extracted as one batch, then box-masked matching; all
    against cv2 (stereo_kat.npz) bit for bit."""
    from visionx_slam_b200.synth import synth_stereo_pair
    g = load_golden("stereo_kat.npz")
    left, right = synth_stereo_pair(1241, 376, 31)
    with vx.VxOrb(n_features=2000, max_width=1241, max_height=376, max_batch=2, max_train=4096) as hx:
        (kl, dl), (kr, dr) = hx.extract_batch([left, right])
        assert_kp_equal(kl, g["kp_left"])
        assert_kp_equal(kr, g["kp_right"])
        assert np.array_equal(dl, g["desc_left"]) and np.array_equal(dr, g["desc_right"])
        qxy, txy = np.stack([kl["x"], kl["y"]], 1), np.stack([kr["x"], kr["y"]], 1)
        knn = hx.knn2_boxed(dl, dr, qxy, txy, g["box"])
        assert _knn_equal(knn, g["knn"])
        m = hx.match_boxed(dl, dr, qxy, txy, g["box"])
        assert np.array_equal(matches_as_rows(m), g["matches"])
        # temporal leg of the same config: square window against the oracle
        mt = hx.match(dl, dr, q_xy=qxy, t_xy=txy, radius=64.0)
        assert np.array_equal(matches_as_rows(mt), matches_as_rows(oracle.match(dl, dr, 0.8, qxy, txy, 64.0)))


@pytest.mark.parametrize("size,n_features,pairs,radius", [((1241, 376), 2000, 4, 48.0), ((640, 480), 1000, 5, 0.0)])
def test_track_stereo_vs_oracle(size, n_features, pairs, radius):
    """BASELINE config 3 through the streaming path: chunks of (left, right) pairs, left -> right through the box
    mask and left_(k-1) -> left_k through the window (or brute force), the last left frame carried across chunks;
    the resident form of the same chunk gives the same bytes."""
    w, hh = size
    frames = synth_stereo_sequence(w, hh, pairs, seed=9)
    box = np.float32([0.0, 64.0, -2.0, 2.0])
    ext = _oracle_extract_many(frames, n_features)
    xy = [np.stack([k["x"], k["y"]], 1) for k, _ in ext]
    with vx.VxOrb(n_features=n_features, max_width=w, max_height=hh, max_batch=2 * pairs, max_train=4096) as hx:
        first = 2 * (pairs // 2)
        f1, m1 = hx.track_stereo(frames[:first], box, radius)           # two chunks: the carry crosses the boundary
        f2, m2 = hx.track_stereo(frames[first:], box, radius)
        got_f, got_m = f1 + f2, m1 + m2
        for i, ((kp, desc), (okp, odesc)) in enumerate(zip(got_f, ext)):
            assert_kp_equal(kp, okp)
            assert np.array_equal(desc, odesc), f"frame {i}"
        for k, (ms, mt) in enumerate(got_m):
            l, r = 2 * k, 2 * k + 1
            om = oracle.match_boxed(ext[l][1], ext[r][1], xy[l], xy[r], box)
            assert np.array_equal(matches_as_rows(ms), matches_as_rows(om)), f"stereo pair {k}"
            if k == 0:
                assert len(mt) == 0
            else:
                p = 2 * (k - 1)
                ot = oracle.match(ext[p][1], ext[l][1], 0.8, xy[p] if radius > 0 else None, xy[l] if radius > 0 else None, radius)
                assert np.array_equal(matches_as_rows(mt), matches_as_rows(ot)), f"temporal pair {k}"
        assert len(got_m[1][0]) > 100 and len(got_m[1][1]) > 100
        # resident form of the whole sequence as one chunk
        hx.track_reset()
        hx.resident_upload(frames)
        hx.resident_run_stereo(box, radius)
        res = hx.resident_fetch()
        for k in range(pairs):
            assert np.array_equal(res[k][2], got_m[k][0]), f"resident stereo pair {k}"
            if k > 0:
                assert np.array_equal(res[pairs + k][2], got_m[k][1]), f"resident temporal pair {k}"


@pytest.mark.parametrize("box", [(0.0, 128.0, -2.0, 2.0), (-5.5, 0.25, -100.0, 3.0), (10.0, 10.0, 0.0, 0.0)])
def test_match_boxed_vs_oracle(h, box):
    rng = np.random.default_rng(int(box[1] * 4) + 7)
    t = rng.integers(0, 256, (700, 32), dtype=np.uint8)
    q = t[rng.permutation(700)[:500]].copy()
    q[:, 3] ^= rng.integers(0, 4, 500).astype(np.uint8)
    qxy = np.round(rng.uniform(0, 300, (500, 2)) * 4).astype(np.float32) / 4  # quarter-pixel lattice: exact ties on the bounds
    txy = np.round(rng.uniform(0, 300, (700, 2)) * 4).astype(np.float32) / 4
    assert _knn_equal(h.knn2_boxed(q, t, qxy, txy, box), oracle.knn2_boxed(q, t, qxy, txy, box))
    assert np.array_equal(matches_as_rows(h.match_boxed(q, t, qxy, txy, box)),
                          matches_as_rows(oracle.match_boxed(q, t, qxy, txy, box)))
