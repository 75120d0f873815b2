import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import visionx_slam_b200 as vx
print(json.dumps({k: v for k, v in bench.bench_per_call(vx, 6531.0).items() if k != "cv2_single_thread"}))
