#!/bin/bash
# A/B of library variants by the bench's resident value: scratch/ab.sh ROUNDS variant.so...  (default library first)
R=${1:-3}; shift
cp visionx_slam_b200/libvxorb.so /tmp/libvxorb_default.so
for i in $(seq $R); do
  for f in /tmp/libvxorb_default.so "$@"; do
    cp $f visionx_slam_b200/libvxorb.so
    timeout 300 python bench.py --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('$f', 'value',round(d['value']),'e2e',round(d['e2e']['value']),'serial_ms',round(d['schedule']['ms_per_step_serial_instrumented'],4))"
  done
done
cp /tmp/libvxorb_default.so visionx_slam_b200/libvxorb.so
