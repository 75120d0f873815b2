"""config 4 (1920x1080, 4000 kp) resident rate at several chunk sizes."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import visionx_slam_b200 as vx
for b in (8, 16, 32):
    out, _, _ = bench.bench_config(vx, 6531.0, 4.5e12, 1920, 1080, 4000, b, 0.0, max(4, 128 // b), f"c4 batch {b}", 32133669)
    print(b, round(out["frames_per_s_resident"]), round(out["extraction_frac_of_hbm_peak"], 4),
          {k: round(v, 4) for k, v in out["stages_ms_per_batch"].items()}, flush=True)
