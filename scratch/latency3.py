"""Where a single-frame Extract call spends its time: per-stage device times at batch 1 (serial, events around
every stage) next to the wall time of the whole call."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import visionx_slam_b200 as vx
from visionx_slam_b200.synth import synth_scene
img = synth_scene(640, 480, 5)
with vx.VxOrb(max_batch=1) as h:
    for _ in range(20): h.extract(img)
    ts = []
    for _ in range(200):
        t0 = time.perf_counter(); h.extract(img); ts.append(time.perf_counter() - t0)
    print("call wall ms (graph replay)", round(float(np.median(ts)) * 1e3, 4))
    h.stage_times_reset(True)
    for _ in range(50): h.extract(img)
    st = h.stage_times()
    print("per-stage device ms per call (serial, instrumented):", {k: round(v[0] / 50, 4) for k, v in st.items()},
          "sum", round(sum(v[0] for v in st.values()) / 50, 4))
    print("launches per call", {k: v[1] / 50 for k, v in st.items()})
