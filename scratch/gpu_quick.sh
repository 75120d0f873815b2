#!/bin/bash
# quick GPU visit: parity tests + bench without the CPU baseline
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
tail -15 gpurun_out/pytest.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_quick.json'))
print('value',round(d['value']),'e2e',round(d['e2e']['value']),'ms',d['ms_per_step'])
for s,e in d['stages'].items(): print(s, round(e['ms_per_step'],4))
PY
