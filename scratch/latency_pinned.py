"""Per-call Extract latency with a pageable and with a pinned source image (what the H2D staging costs)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
import visionx_slam_b200 as vx
from visionx_slam_b200.synth import synth_scene
a = synth_scene(640, 480, 11)
with vx.VxOrb(max_batch=1) as h:
    kp = np.zeros(h.kp_cap, vx.KEYPOINT_DTYPE); desc = np.zeros((h.kp_cap, 32), np.uint8)
    pin = vx.PinnedArray((480, 640), np.uint8); pin.array[:] = a
    pk = vx.PinnedArray((h.kp_cap,), vx.KEYPOINT_DTYPE); pd = vx.PinnedArray((h.kp_cap, 32), np.uint8)
    print("pageable in, pageable out", bench.median_ms(lambda: h.extract_into(a, kp, desc), 300))
    print("pinned in,   pageable out", bench.median_ms(lambda: h.extract_into(pin.array, kp, desc), 300))
    print("pinned in,   pinned out  ", bench.median_ms(lambda: h.extract_into(pin.array, pk.array, pd.array), 300))
    h.stage_times_reset(True)
    for _ in range(50): h.extract_into(pin.array, kp, desc)
    st = h.stage_times(); h.stage_times_reset(False)
    print("device stage ms per call (instrumented, no graph):", {k: round(v[0] / 50, 4) for k, v in st.items()}, "sum", round(sum(v[0] for v in st.values()) / 50, 4))
