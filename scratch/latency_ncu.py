"""Kernel durations of ONE single-frame Extract call (run under ncu --metrics gpu__time_duration.sum)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import visionx_slam_b200 as vx
from visionx_slam_b200.synth import synth_scene
os.environ["VXORB_NO_GRAPH"] = "1"
img = synth_scene(640, 480, 5)
with vx.VxOrb(max_batch=1) as h:
    for _ in range(3): h.extract(img)
