#!/bin/bash
# ncu --set full capture of one launch of the kernels matching $1 (regex), after warm-up; report name $2
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name "regex:$1" --launch-skip ${3:-3} -c ${4:-1} -o gpurun_out/$2 -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_$2.log 2>&1
tail -3 gpurun_out/ncu_$2.log
