"""How much of the e2e rate does the D2H side cost?  Same streaming loop as bench.py with fewer outputs."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import visionx_slam_b200 as vx
from visionx_slam_b200.synth import synth_sequence
B, W, H = 128, 640, 480
base = synth_sequence(W, H, 32, seed=2026)
seq = np.concatenate([base if r % 2 == 0 else base[::-1] for r in range(4)])[:B]
h = vx.VxOrb(max_width=W, max_height=H, max_batch=B)
cap = h.kp_cap
pin = [vx.PinnedArray((B, H, W), np.uint8) for _ in range(3)]
for p in pin: p.array[:] = seq
outs = [dict(kp=vx.PinnedArray((B, cap), vx.KEYPOINT_DTYPE), desc=vx.PinnedArray((B, cap, 32), np.uint8), nk=vx.PinnedArray((B,), np.int32),
             mt=vx.PinnedArray((B, cap), vx.MATCH_DTYPE), nm=vx.PinnedArray((B,), np.int32)) for _ in range(3)]
def run(n, mode):
    infl = []
    for i in range(n):
        k = i % 3
        if len(infl) == 3: h.track_wait(infl.pop(0))
        o = outs[k]
        if mode == "all": a = (o["kp"].array, o["desc"].array, o["nk"].array, o["mt"].array, o["nm"].array)
        elif mode == "nodesc": a = (o["kp"].array, None, o["nk"].array, o["mt"].array, o["nm"].array)
        else: a = (None, None, o["nk"].array, None, o["nm"].array)
        infl.append(h.track_submit(pin[k].array, 32.0, *a))
    for s in infl: h.track_wait(s)
for mode in ("all", "nodesc", "counts", "all"):
    h.track_reset(); run(6, mode); h.sync()
    t0 = time.perf_counter(); run(40, mode); dt = time.perf_counter() - t0
    print(mode, round(40 * B / dt), "frames/s")
