"""Single-frame latency of the drop-in calls (what Tracking::ProcessFrame pays per frame):
vxorb_extract on one 640x480 frame and vxorb_match on 1000x1000 descriptors, host buffers in and out."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import visionx_slam_b200 as vx
from visionx_slam_b200.synth import synth_sequence
fr = synth_sequence(640, 480, 8, seed=4)
with vx.VxOrb(max_batch=1) as h:
    outs = [h.extract(f) for f in fr]
    for name, fn in [("extract", lambda i: h.extract(fr[i % 8])),
                     ("extract_info", lambda i: h.extract_info(fr[i % 8])),
                     ("match", lambda i: h.match(outs[i % 7][1], outs[i % 7 + 1][1])),
                     ("match_filtered", lambda i: h.match_filtered(outs[i % 7][1], outs[i % 7 + 1][1]))]:
        for i in range(20): fn(i)
        ts = []
        for i in range(200):
            t0 = time.perf_counter(); fn(i); ts.append(time.perf_counter() - t0)
        ts = np.array(ts) * 1e3
        print(f"{name}: median {np.median(ts):.3f} ms, p10 {np.percentile(ts,10):.3f}, p90 {np.percentile(ts,90):.3f}")
rng = np.random.default_rng(0)
q4, t16 = rng.integers(0, 256, (4000, 32), dtype=np.uint8), rng.integers(0, 256, (16384, 32), dtype=np.uint8)
with vx.VxOrb(max_batch=1, max_train=20000) as h:
    for i in range(5): h.match(q4, t16)
    ts = []
    for i in range(30):
        t0 = time.perf_counter(); h.match(q4, t16); ts.append(time.perf_counter() - t0)
    print(f"match 4000x16384: median {np.median(ts)*1e3:.3f} ms  (popc floor 0.116 ms)")
try:
    import cv2
    cv2.setNumThreads(1)
    orb = cv2.ORB_create(1000, 1.2, 8); bf = cv2.BFMatcher(cv2.NORM_HAMMING)
    ts = []
    for i in range(30):
        t0 = time.perf_counter(); orb.detectAndCompute(fr[i % 8], None); ts.append(time.perf_counter() - t0)
    print(f"cv2 detectAndCompute 1 thread: median {np.median(ts)*1e3:.2f} ms")
    ts = []
    for i in range(30):
        t0 = time.perf_counter(); bf.knnMatch(outs[i % 7][1], outs[i % 7 + 1][1], 2); ts.append(time.perf_counter() - t0)
    print(f"cv2 knnMatch 1 thread: median {np.median(ts)*1e3:.2f} ms")
except ImportError:
    pass
