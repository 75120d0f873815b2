"""Randomised parity soak: odd frame sizes, parameters and descriptor-set sizes against the CPU oracle."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import oracle
import visionx_slam_b200 as vx
from visionx_slam_b200.synth import synth_scene

rng = np.random.default_rng(int(sys.argv[1]) if len(sys.argv) > 1 else 1)
bad = 0
def kp_eq(a, b):
    return len(a) == len(b) and all(np.array_equal(a[k], b[k]) for k in ("x", "y", "response", "angle", "octave"))
for case in range(int(sys.argv[2]) if len(sys.argv) > 2 else 40):
    w, hh = int(rng.integers(40, 900)), int(rng.integers(40, 700))
    nf = int(rng.choice([50, 300, 1000, 2500])); nl = int(rng.integers(1, 10)); sf = float(rng.choice([1.1, 1.2, 1.3, 1.5, 2.0]))
    kind = rng.choice(["scene", "noise", "grad"])
    if kind == "scene": img = synth_scene(w, hh, int(rng.integers(0, 1000)))
    elif kind == "noise": img = rng.integers(0, 256, (hh, w), dtype=np.uint8)
    else: img = ((np.add.outer(np.arange(hh), np.arange(w)) * 3) % 256).astype(np.uint8)
    b = int(rng.choice([1, 2, 3, 8, 9]))  # 8, 9: the batched (throughput) kernel instantiations
    try:
        with vx.VxOrb(n_features=nf, scale_factor=sf, n_levels=nl, max_width=w, max_height=hh, max_batch=b,
                      max_keypoints=nf * 3 + 512) as h:
            frames = [img] + [np.roll(img, int(rng.integers(1, 9)), axis=1) for _ in range(b - 1)]
            outs = h.extract_batch(frames) if b > 1 else [h.extract(img)]
            for f, (kp, desc) in zip(frames, outs):
                okp, odesc = oracle.orb_extract(f, nf, sf, nl)
                if not (kp_eq(kp, okp) and np.array_equal(desc, odesc)):
                    bad += 1
                    print("EXTRACT MISMATCH", case, w, hh, nf, nl, sf, kind, b, len(kp), len(okp))
            if b > 1:
                tr = h.track(np.stack(frames), 24.0)
                for f, (kp, desc, m) in zip(frames, tr):
                    okp, odesc = oracle.orb_extract(f, nf, sf, nl)
                    if not (kp_eq(kp, okp) and np.array_equal(desc, odesc)):
                        bad += 1
                        print("TRACK MISMATCH", case, w, hh, nf, nl, sf, kind, b)
    except vx.VxOrbError as e:
        print("ERROR", case, w, hh, nf, nl, sf, kind, b, e)
        bad += 1
with vx.VxOrb(max_train=9000) as h:
    for case in range(30):
        nq, nt = int(rng.integers(1, 5000)), int(rng.integers(1, 8000))
        t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
        q = t[rng.integers(0, nt, nq)].copy(); q[:, int(rng.integers(0, 32))] ^= rng.integers(0, 16, nq).astype(np.uint8)
        qxy = (rng.uniform(0, 500, (nq, 2))).astype(np.float32); txy = (rng.uniform(0, 500, (nt, 2))).astype(np.float32)
        mode = case % 3
        if mode == 0:
            a, o = h.knn2(q, t), oracle.knn2(q, t)
            if nq <= vx.STATIC_MAX_QUERY and nt <= vx.STATIC_MAX_TRAIN:  # the allocation-free entry point, same rows
                st = vx.knn2_static(q, t)
                if not (np.array_equal(st["idx"], o["idx"]) and np.array_equal(np.where(st["idx"] >= 0, st["dist"], 0), np.where(o["idx"] >= 0, o["dist"], 0))):
                    bad += 1
                    print("STATIC MISMATCH", case, nq, nt)
        elif mode == 1:
            r = float(rng.choice([8, 32, 100])); a, o = h.knn2(q, t, qxy, txy, r), oracle.knn2(q, t, qxy, txy, r)
        else:
            box = [-float(rng.uniform(0, 50)), float(rng.uniform(0, 80)), -3.0, float(rng.uniform(0, 30))]
            a, o = h.knn2_boxed(q, t, qxy, txy, box), oracle.knn2_boxed(q, t, qxy, txy, box)
        ok = np.array_equal(a["idx"], o["idx"]) and np.array_equal(np.where(a["idx"] >= 0, a["dist"], 0), np.where(o["idx"] >= 0, o["dist"], 0))
        if not ok:
            bad += 1
            print("MATCH MISMATCH", case, nq, nt, mode)
print("soak done, mismatches:", bad)
