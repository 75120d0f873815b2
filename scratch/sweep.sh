#!/bin/bash
# runs parity (FAST + C1 extract) and the bench for every library variant under scratch/variants
cp visionx_slam_b200/libvxorb.so /tmp/libvxorb_default.so
for f in /tmp/libvxorb_default.so scratch/variants/*.so; do
  cp $f visionx_slam_b200/libvxorb.so
  echo "== $f"
  timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "fast or blur or resize or pyramid or extract_c1 or extract_large or adversarial" 2>&1 | tail -1
  timeout 300 python bench.py --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('value',round(d['value']),'e2e',round(d['e2e']['value']),' '.join(f\"{s}={e['ms_per_step']:.4f}\" for s,e in d['stages'].items()))"
done
cp /tmp/libvxorb_default.so visionx_slam_b200/libvxorb.so
