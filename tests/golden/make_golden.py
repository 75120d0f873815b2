"""Generate the golden vectors under tests/golden/ from OpenCV itself (cv2 4.13.0).

The reference (VisionX-SLAM) delegates the whole hot path to OpenCV:
  core/feature/orb_extractor.cpp:6,13   cv::ORB::create(1000,1.2f,8)->detectAndCompute
  core/feature/orb_matcher.cpp:22-36    cv::BFMatcher(NORM_HAMMING).knnMatch(q,t,knn,2) + ratio
cv2 exposes exactly those entry points, so these vectors are outputs of the reference's own
arithmetic.  Run in the build container (cv2 does not have to exist on the GPU box):

    python tests/golden/make_golden.py

Keypoints/descriptors are stored in the canonical order (octave, y, x) (cv::ORB's own order is an
artefact of std::nth_element) and the matcher is run on the re-ordered rows, so
indices are comparable bit for bit.
"""
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from visionx_slam_b200.synth import sha256, synth_scene, synth_stereo_pair  # noqa: E402

KP = np.dtype([("x", "<f4"), ("y", "<f4"), ("response", "<f4"), ("angle", "<f4"), ("octave", "<i4")])


def level_scales(n_levels, sf=np.float32(1.2)):
    return [np.float32(pow(float(sf), l)) for l in range(n_levels)]


def cv_extract(img, n_features, n_levels=8, sf=1.2):
    orb = cv2.ORB_create(n_features, sf, n_levels)
    kps, desc = orb.detectAndCompute(img, None)
    if len(kps) == 0:
        return np.zeros(0, KP), np.zeros((0, 32), np.uint8)
    sc = level_scales(n_levels)
    rows = []
    for i, k in enumerate(kps):
        s = sc[k.octave]
        rows.append((k.octave, int(round(k.pt[1] / s)), int(round(k.pt[0] / s)), i))
    rows.sort()
    out = np.zeros(len(kps), KP)
    for j, (_, _, _, i) in enumerate(rows):
        k = kps[i]
        out[j] = (k.pt[0], k.pt[1], k.response, k.angle, k.octave)
    return out, np.ascontiguousarray(desc[[r[3] for r in rows]])


def cv_knn2(q, t, mask=None):
    out = np.zeros(len(q), np.dtype([("idx", "<i4", (2,)), ("dist", "<i4", (2,))]))
    out["idx"] = -1
    if len(q) == 0 or len(t) == 0:
        return out
    bf = cv2.BFMatcher(cv2.NORM_HAMMING)
    knn = bf.knnMatch(q, t, 2) if mask is None else bf.knnMatch(q, t, 2, mask=mask)
    for i, row in enumerate(knn):
        for k, m in enumerate(row[:2]):
            assert m.queryIdx == i
            out[i]["idx"][k] = m.trainIdx
            out[i]["dist"][k] = int(m.distance)
    return out


def ratio_matches(knn, nn_ratio=np.float32(0.8)):
    rows = []
    for i, r in enumerate(knn):
        if r["idx"][1] < 0:
            continue
        if np.float32(r["dist"][0]) < nn_ratio * np.float32(r["dist"][1]):
            rows.append((i, r["idx"][0], r["dist"][0]))
    return np.array(rows, np.int32).reshape(-1, 3)


def cv_crosscheck(q, t):
    ms = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=True).match(q, t)
    rows = sorted((m.queryIdx, m.trainIdx, int(m.distance)) for m in ms)
    return np.array(rows, np.int32).reshape(-1, 3)


def window_mask(qxy, txy, radius):
    dx = np.abs(qxy[:, None, 0] - txy[None, :, 0]) <= np.float32(radius)
    dy = np.abs(qxy[:, None, 1] - txy[None, :, 1]) <= np.float32(radius)
    return (dx & dy).astype(np.uint8)


def stage_vectors(img, n_levels=8):
    """Per-stage cv2 primitives on one frame."""
    out = {}
    h, w = img.shape
    sc = level_scales(n_levels)
    lvl = img
    k32 = cv2.getGaussianKernel(7, 2, cv2.CV_32F)
    fast = cv2.FastFeatureDetector_create(20, True)
    for l in range(n_levels):
        if l > 0:
            inv = np.float32(1.0) / sc[l]
            dw = int(np.rint(np.float32(w) * inv))
            dh = int(np.rint(np.float32(h) * inv))
            lvl = cv2.resize(lvl, (dw, dh), interpolation=cv2.INTER_LINEAR_EXACT)
        out[f"level{l}_shape"] = np.array(lvl.shape, np.int32)
        out[f"level{l}_sha"] = np.array(sha256(lvl))
        bl = cv2.sepFilter2D(lvl, cv2.CV_8U, k32, k32, borderType=cv2.BORDER_REFLECT_101)
        out[f"blur{l}_sha"] = np.array(sha256(bl))
        kps = fast.detect(lvl)
        out[f"fast{l}"] = np.array([(int(k.pt[0]), int(k.pt[1]), int(k.response)) for k in kps],
                                   np.int16).reshape(-1, 3)
    return out


def grid_image(w=640, h=480):
    img = np.zeros((h, w), np.uint8)
    for gy in range(12):
        for gx in range(12):
            x0, y0 = 40 + gx * 48, 40 + gy * 34
            img[y0:y0 + 17, x0:x0 + 24] = 255
    return img


def main():
    # ---- C1: 640x480 pair, 1000 kp (BASELINE configs[0]) -------------------------------------
    a = synth_scene(640, 480, 11)
    b = np.roll(a, (3, 5), axis=(0, 1))
    ka, da = cv_extract(a, 1000)
    kb, db = cv_extract(b, 1000)
    knn = cv_knn2(da, db)
    qxy = np.stack([ka["x"], ka["y"]], 1)
    txy = np.stack([kb["x"], kb["y"]], 1)
    knn_w = cv_knn2(da, db, window_mask(qxy, txy, 32.0))
    g = dict(frame_a=a, frame_b=b, kp_a=ka, desc_a=da, kp_b=kb, desc_b=db, knn=knn,
             matches=ratio_matches(knn), knn_win32=knn_w, matches_win32=ratio_matches(knn_w),
             crosscheck=cv_crosscheck(da, db))
    g.update({"a_" + k: v for k, v in stage_vectors(a).items()})
    np.savez_compressed(os.path.join(HERE, "c1_640x480.npz"), **g)
    print("c1", len(ka), len(kb), len(g["matches"]), len(g["matches_win32"]), len(g["crosscheck"]))

    # ---- C3 / C4 shapes: frames are regenerated from the seed, pinned by sha256 ---------------
    for name, (w, h, n, seed) in {"c3_1241x376": (1241, 376, 2000, 7), "c4_1920x1080": (1920, 1080, 4000, 3)}.items():
        img = synth_scene(w, h, seed)
        k, d = cv_extract(img, n)
        st = stage_vectors(img)
        keep = {kk: v for kk, v in st.items() if not kk.startswith("fast") or kk in ("fast0", "fast7")}
        np.savez_compressed(os.path.join(HERE, name + ".npz"), seed=seed, w=w, h=h, n_features=n,
                            frame_sha=np.array(sha256(img)), kp=k, desc=d, **keep)
        print(name, len(k))

    # ---- adversarial extraction cases ----------------------------------------------------------
    rng = np.random.default_rng(5)
    adv = {}
    cases = {
        "grid": grid_image(),                                     # ties kept: more kps than quota
        "noise": rng.integers(0, 256, (240, 320), dtype=np.uint8),  # dense corners, heavy score ties
        "blank": np.full((480, 640), 128, np.uint8),
        "tiny60": rng.integers(0, 256, (60, 60), dtype=np.uint8),   # every level <= 62 px: 0 kp
        "small80x70": rng.integers(0, 256, (70, 80), dtype=np.uint8),
        "lowtex": (synth_scene(640, 480, 9) // 8 + 100).astype(np.uint8),
        "odd_357x291": synth_scene(357, 291, 2),
    }
    for name, img in cases.items():
        k, d = cv_extract(img, 1000)
        adv[name + "_img"] = img
        adv[name + "_kp"] = k
        adv[name + "_desc"] = d
        print("adv", name, len(k))
    # other parameter sets through the same code
    img = synth_scene(640, 480, 8)
    for tag, (n, sf, nl) in {"p500_4": (500, 1.2, 4), "p300_s15_5": (300, 1.5, 5), "p2000_8": (2000, 1.2, 8)}.items():
        k, d = cv_extract(img, n, nl, sf)
        adv[tag + "_kp"] = k
        adv[tag + "_desc"] = d
        adv[tag + "_params"] = np.array([n, sf, nl], np.float64)
        print("adv", tag, len(k))
    adv["params_img_seed"] = np.array(8)
    # BGR input (system.cpp:46): ORB on a colour frame
    bgr = np.stack([synth_scene(320, 240, s) for s in (21, 22, 23)], axis=2)
    k, d = cv_extract(bgr, 500)
    adv["bgr_img"] = bgr
    adv["bgr_gray"] = cv2.cvtColor(bgr, cv2.COLOR_BGR2GRAY)
    adv["bgr_kp"] = k
    adv["bgr_desc"] = d
    np.savez_compressed(os.path.join(HERE, "adversarial.npz"), **adv)

    # ---- matcher known-answer tests ------------------------------------------------------------
    m = {}
    rng = np.random.default_rng(17)
    # heavy ties: descriptors drawn from a tiny alphabet, with duplicates
    base = rng.integers(0, 256, (6, 32), dtype=np.uint8)
    q = base[rng.integers(0, 6, 200)]
    t = base[rng.integers(0, 6, 150)]
    flip = rng.integers(0, 150, 40)
    t = t.copy()
    t[flip, 0] ^= 1
    for tag, (qq, tt) in {"ties": (q, t), "nt1": (q[:50], t[:1]), "nt2": (q[:50], t[:2]),
                          "rand": (rng.integers(0, 256, (333, 32), dtype=np.uint8),
                                   rng.integers(0, 256, (517, 32), dtype=np.uint8))}.items():
        kn = cv_knn2(qq, tt)
        m[tag + "_q"], m[tag + "_t"], m[tag + "_knn"] = qq, tt, kn
        m[tag + "_matches"] = ratio_matches(kn)
        m[tag + "_crosscheck"] = cv_crosscheck(qq, tt)
    # near-duplicate structure so the ratio test passes often
    t2 = rng.integers(0, 256, (400, 32), dtype=np.uint8)
    q2 = t2[rng.permutation(400)[:300]].copy()
    noise_bits = rng.integers(0, 256, (300, 4))
    for i in range(300):
        for bpos in noise_bits[i]:
            q2[i, bpos // 8] ^= 1 << (bpos % 8)
    kn = cv_knn2(q2, t2)
    m["dup_q"], m["dup_t"], m["dup_knn"], m["dup_matches"] = q2, t2, kn, ratio_matches(kn)
    m["dup_crosscheck"] = cv_crosscheck(q2, t2)
    qxy = rng.uniform(0, 640, (300, 2)).astype(np.float32)
    txy = rng.uniform(0, 640, (400, 2)).astype(np.float32)
    knw = cv_knn2(q2, t2, window_mask(qxy, txy, 64.0))
    m["dup_qxy"], m["dup_txy"], m["dup_knn_win64"], m["dup_matches_win64"] = qxy, txy, knw, ratio_matches(knw)
    np.savez_compressed(os.path.join(HERE, "matcher_kat.npz"), **m)
    print("matcher", {k: len(v) for k, v in m.items() if k.endswith("matches")})


def caller_kat():
    """cv::meanStdDev on the gray frame exactly as Tracking::CheckImageQuality calls it
    (core/frontend/tracking.cpp:120-139: cvtColor(BGR2GRAY) then meanStdDev) -> caller_kat.npz."""
    rng = np.random.default_rng(23)
    imgs = {
        "synth": synth_scene(640, 480, 11),
        "noise_odd": rng.integers(0, 256, (61, 97), dtype=np.uint8),
        "const": np.full((100, 100), 7, np.uint8),
        "bright": rng.integers(250, 256, (131, 203), dtype=np.uint8),
        "dark_w3": rng.integers(0, 9, (75, 3), dtype=np.uint8),
        "bgr": rng.integers(0, 256, (77, 123, 3), dtype=np.uint8),
    }
    out = {}
    for k, im in imgs.items():
        g = cv2.cvtColor(im, cv2.COLOR_BGR2GRAY) if im.ndim == 3 else im
        mean, sd = cv2.meanStdDev(g)
        out[k + "_img"] = im
        out[k + "_stats"] = np.array([mean[0, 0], sd[0, 0]], np.float64)
    np.savez_compressed(os.path.join(HERE, "caller_kat.npz"), **out)
    print("caller", {k: v.tolist() for k, v in out.items() if k.endswith("_stats")})


def box_mask(qxy, txy, box):
    dx = qxy[:, None, 0] - txy[None, :, 0]
    dy = qxy[:, None, 1] - txy[None, :, 1]
    b = np.asarray(box, np.float32)
    return ((dx >= b[0]) & (dx <= b[1]) & (dy >= b[2]) & (dy <= b[3])).astype(np.uint8)


def stereo_kat():
    """BASELINE config 3 shape: KITTI-size 1241x376 rectified pair, 2000 kp per image, left -> right
    matching through cv knnMatch with the epipolar-band mask 0 <= xl - xr <= 128, |yl - yr| <= 2
    -> stereo_kat.npz.  Frames come from synth_stereo_pair (only their hashes are stored)."""
    left, right = synth_stereo_pair(1241, 376, 31)
    kl, dl = cv_extract(left, 2000)
    kr, dr = cv_extract(right, 2000)
    qxy = np.stack([kl["x"], kl["y"]], 1)
    txy = np.stack([kr["x"], kr["y"]], 1)
    box = np.array([0.0, 128.0, -2.0, 2.0], np.float32)
    knn = cv_knn2(dl, dr, box_mask(qxy, txy, box))
    m = ratio_matches(knn)
    np.savez_compressed(os.path.join(HERE, "stereo_kat.npz"), left_sha=sha256(left), right_sha=sha256(right), box=box,
                        kp_left=kl, desc_left=dl, kp_right=kr, desc_right=dr, knn=knn, matches=m)
    print("stereo", len(kl), len(kr), "matches", len(m))


def masked_kat():
    """cv knnMatch(k=2, mask) with an arbitrary nq x nt CV_8U mask (SURVEY.md 8b vxorb_match_masked): few distinct
    descriptors (heavy distance ties -> lowest allowed train index), rows with no / one / all pairs allowed
    -> masked_kat.npz."""
    rng = np.random.default_rng(77)
    base = rng.integers(0, 256, (12, 32), dtype=np.uint8)
    t = base[rng.integers(0, 12, 700)].copy()
    q = base[rng.integers(0, 12, 300)].copy()
    t[::4, 5] ^= 0x33
    q[::3, 9] ^= 0x0c
    t[600:] = rng.integers(0, 256, (100, 32), dtype=np.uint8)      # distinct rows with near-copies among the queries
    q[200:] = t[600:]
    q[200:, 3] ^= 0x41
    mask = (rng.random((300, 700)) < 0.15).astype(np.uint8)
    mask[np.arange(200, 300), np.arange(600, 700)] = (np.arange(100) % 3 != 0)  # a third of the true partners masked out
    mask[5] = 0                      # no neighbour at all
    mask[6] = 0
    mask[6, 123] = 1                 # exactly one neighbour -> knn row of size 1 -> never a ratio match
    mask[7] = 255                    # everything allowed, any non-zero value counts
    knn = cv_knn2(q, t, mask)
    m = ratio_matches(knn)
    np.savez_compressed(os.path.join(HERE, "masked_kat.npz"), q=q, t=t, mask=mask, knn=knn, matches=m)
    print("masked", len(q), len(t), "allowed", int(mask.astype(bool).sum()), "matches", len(m))


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if which in ("all", "masked"):
        masked_kat()
    if which in ("all",):
        main()
    if which in ("all", "caller"):
        caller_kat()
    if which in ("all", "stereo"):
        stereo_kat()
