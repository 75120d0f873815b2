import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import visionx_slam_b200 as vx
from visionx_slam_b200.synth import synth_scene
def t(fn, n=200):
    for i in range(20): fn()
    ts=[]
    for i in range(n):
        t0=time.perf_counter(); fn(); ts.append(time.perf_counter()-t0)
    return np.median(ts)*1e3
with vx.VxOrb(max_batch=1) as h:
    img=synth_scene(640,480,5); blank=np.zeros((480,640),np.uint8); small=synth_scene(128,96,5)
    print("640x480 scene", round(t(lambda: h.extract(img)),4))
    print("640x480 blank", round(t(lambda: h.extract(blank)),4))
with vx.VxOrb(max_batch=1, max_width=128, max_height=96) as h:
    print("128x96 scene ", round(t(lambda: h.extract(small)),4))
