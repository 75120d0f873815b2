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
    r = [bench.median_ms(lambda: h.extract_into(pin.array, kp, desc), 400) for _ in range(3)]
    print(os.environ.get("VXORB_LIB", "default")[-20:], "NO_GRAPH" if os.environ.get("VXORB_NO_GRAPH") else "graph", [round(x[0], 4) for x in r], "min", round(min(x[1] for x in r), 4))
