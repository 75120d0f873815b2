"""per-stage device time of single-frame extraction (instrumented, serial) and the wall time of the graph path."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import visionx_slam_b200 as vx
from visionx_slam_b200.synth import synth_scene
img = synth_scene(640, 480, 11)
with vx.VxOrb() as h:
    kp = np.zeros(h.kp_cap, vx.KEYPOINT_DTYPE); desc = np.zeros((h.kp_cap, 32), np.uint8)
    for _ in range(20): h.extract_into(img, kp, desc)
    t = []
    for _ in range(300):
        t0 = time.perf_counter(); h.extract_into(img, kp, desc); t.append(time.perf_counter() - t0)
    print("graph path wall ms median", round(float(np.median(t)) * 1e3, 4), "min", round(min(t) * 1e3, 4))
    h.stage_times_reset(True)
    for _ in range(50): h.extract_into(img, kp, desc)
    st = h.stage_times()
    print({k: (round(v[0] / 50 * 1e3, 2), v[1] // 50) for k, v in st.items()}, "us per call (stage, launches)")
    h.stage_times_reset(False)
