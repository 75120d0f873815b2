#!/usr/bin/env python
"""One serial resident pass of the benchmark's 128-frame chunk per matcher, for ncu.

    python profiles/prof_step.py [batch]                       # plain run (must exit 0 before the ncu run)
    ncu --set full --clock-control none --import-source on -s 42 -c 26 -o gpurun_out/r2x python profiles/prof_step.py

Launch order after the warm-up (3 unpack launches of the upload + 3 passes x 13 launches = 42 to skip):
pass A (brute-force matcher, radius 0): 7 x resize_band, fast, select, blur, describe, knn2<1>, match_compact;
pass B (windowed matcher, radius 32): the same with knn2_window.  Stage instrumentation is on, so the stages run
one after another on one stream (the production schedule overlaps the blur with FAST + select)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import visionx_slam_b200 as vx  # noqa: E402
from visionx_slam_b200.synth import synth_sequence  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
base = synth_sequence(640, 480, min(B, 32), seed=2026)
reps = (B + len(base) - 1) // len(base)
seq = np.concatenate([base if r % 2 == 0 else base[::-1] for r in range(reps)])[:B]
with vx.VxOrb(max_batch=B) as h:
    h.resident_upload(seq)
    h.stage_times_reset(True)
    for _ in range(3):
        h.resident_run(0.0)
    h.sync()
    h.resident_run(0.0)
    h.resident_run(32.0)
    h.sync()
    st = h.stage_times()
    print({k: round(v[0] / 5, 4) for k, v in st.items()}, "launches", h.launch_count())
