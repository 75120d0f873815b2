#!/usr/bin/env python
"""Stage times (serial instrumented pass, ms per 128-frame chunk) and the pipelined resident rate of every library
variant under visionx_slam_b200/variants/ plus the default build; one subprocess per variant.
    python scratch/ab_run.py [radius] [reps]"""
import glob
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import sys, json, numpy as np
sys.path.insert(0, %r)
import visionx_slam_b200 as vx
from visionx_slam_b200.synth import synth_sequence
radius, reps = float(sys.argv[1]), int(sys.argv[2])
B = 128
base = synth_sequence(640, 480, 32, seed=2026)
seq = np.concatenate([base if r %% 2 == 0 else base[::-1] for r in range(4)])[:B]
with vx.VxOrb(max_batch=B) as h:
    h.resident_upload(seq)
    for _ in range(6): h.resident_run(radius)
    h.sync()
    best = 1e9
    for _ in range(3):
        h.timer_begin()
        for _ in range(reps): h.resident_run(radius)
        best = min(best, h.timer_end() / reps)
    h.stage_times_reset(True)
    for _ in range(reps): h.resident_run(radius)
    st = h.stage_times()
    h.stage_times_reset(False)
    res = h.resident_fetch()
    import hashlib
    dig = hashlib.sha256(b"".join(r[0].tobytes() + r[1].tobytes() + r[2].tobytes() for r in res)).hexdigest()[:12]
    print(json.dumps({"pipelined_ms": best, "stages": {k: v[0] / reps for k, v in st.items()}, "digest": dig}))
''' % ROOT

radius = sys.argv[1] if len(sys.argv) > 1 else "0"
reps = sys.argv[2] if len(sys.argv) > 2 else "20"
libs = [("default", None)] + [(os.path.basename(p)[9:-3], p) for p in sorted(glob.glob(os.path.join(ROOT, "visionx_slam_b200", "variants", "libvxorb_*.so")))]
for name, path in libs + libs[:1]:  # the default build again at the end: run-to-run spread
    env = dict(os.environ)
    if path:
        env["VXORB_LIB"] = path
    r = subprocess.run([sys.executable, "-c", CHILD, radius, reps], env=env, capture_output=True, text=True)
    if r.returncode != 0:
        print(name, "FAILED", r.stderr[-400:])
        continue
    d = json.loads(r.stdout.strip().splitlines()[-1])
    st = d["stages"]
    print(f"{name:14s} pipelined {d['pipelined_ms']:.4f} ms/chunk ({128 / d['pipelined_ms']:.1f} k frames/s)  " +
          " ".join(f"{k} {v:.4f}" for k, v in st.items()) + f"  sum {sum(st.values()):.4f}  digest {d['digest']}", flush=True)
