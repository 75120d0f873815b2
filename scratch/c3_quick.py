import sys, json; sys.path.insert(0,__import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.abspath(__file__))))
import bench, visionx_slam_b200 as vx
with vx.VxOrb() as h: popc = h.measure_popc_rate()
print(json.dumps(bench.bench_c3(vx, 6531.0, popc)))
