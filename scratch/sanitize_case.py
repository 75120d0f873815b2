#!/usr/bin/env python
"""Small run that touches every kernel family in both batch shapes (batch 8 = the throughput instantiations,
batch 1 = the latency shapes / CUDA graph), for compute-sanitizer:
    compute-sanitizer --tool memcheck  python scratch/sanitize_case.py
    compute-sanitizer --tool racecheck python scratch/sanitize_case.py
Results are checked against the oracle so that a tool-induced slowdown cannot hide a wrong answer."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle  # noqa: E402
import visionx_slam_b200 as vx  # noqa: E402
from visionx_slam_b200.synth import synth_scene, synth_sequence  # noqa: E402

frames = synth_sequence(640, 480, 8, seed=3)
ext = [oracle.orb_extract(f) for f in frames]
with vx.VxOrb(max_batch=8) as h:
    for radius in (32.0, 0.0):
        h.track_reset()
        got = h.track(frames, radius)
        for i, ((kp, desc, m), (okp, odesc)) in enumerate(zip(got, ext)):
            assert np.array_equal(kp, okp) and np.array_equal(desc, odesc), f"frame {i}"
            if i:
                pk, pd = ext[i - 1]
                win = radius > 0
                om = oracle.match(pd, odesc, 0.8, np.stack([pk["x"], pk["y"]], 1) if win else None,
                                  np.stack([okp["x"], okp["y"]], 1) if win else None, radius)
                assert np.array_equal(m, om), f"matches frame {i} radius {radius}"
    ka, da, ia = h.extract_id(frames[0])
    kb, db, ib, info = h.extract_id(frames[1], with_info=True)
    assert np.array_equal(h.match_ids(ia, ib), oracle.match(da, db, 0.8))
    assert np.array_equal(h.match_crosscheck(da, db), oracle.match_crosscheck(da, db))
    assert np.array_equal(h.match_filtered(da, db), oracle.postfilter(oracle.match(da, db, 0.8)))
    qxy, txy = np.stack([ka["x"], ka["y"]], 1), np.stack([kb["x"], kb["y"]], 1)
    box = np.float32([-40, 40, -3, 3])
    assert np.array_equal(h.match_boxed(da, db, qxy, txy, box), oracle.match_boxed(da, db, qxy, txy, box))
    small = synth_scene(413, 297, 5)
    kp, desc = h.extract(small)
    okp, odesc = oracle.orb_extract(small)
    assert np.array_equal(kp, okp) and np.array_equal(desc, odesc)
    img = synth_scene(200, 150, 9)
    assert np.array_equal(h.debug_blur(img), oracle.blur7(img))
    c = h.debug_fast(img, 20, 3)
    assert len(c) == len(oracle.fast9_nms(img, 20, 3))
    for rows in (16, 8, 4):
        assert np.array_equal(h.debug_resize_band(img, 167, 125, rows), oracle.resize_linear_exact(img, 167, 125))
print("sanitize case ok: every result equals the oracle's")
