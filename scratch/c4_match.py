"""config-4-shaped temporal matching (4000 x 4000 brute force) at several pair counts, per library variant."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import visionx_slam_b200 as vx
for b in (2, 4, 8, 16, 32):
    out, _, _ = bench.bench_config(vx, 6531.0, 4.5e12, 1920, 1080, 4000, b, 0.0, max(4, 64 // b), f"c4 batch {b}", 32133669)
    print(os.environ.get("VXORB_LIB", "default")[-16:], b, round(out["frames_per_s_resident"]), "match ms/pair", round(out["match"]["ms_per_batch"] / b, 4), flush=True)
