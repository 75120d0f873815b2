#!/bin/bash
# One gpurun call: gpu tests, bench line, ncu launch list of the bench command, ncu full set of one serial step.
# usage: scratch/gpu_round.sh <tag> [skip_tests]
set -u
TAG=${1:-r2a}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv -lms 500 > $OUT/${TAG}_clocks.csv 2>/dev/null &
SMI=$!
if [ "${2:-}" != "skip_tests" ]; then
  timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/${TAG}_pytest.log
  tail -5 $OUT/${TAG}_pytest.log
fi
timeout 600 python bench.py --steps 20 --warmup 3 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"
tail -c 600 $OUT/${TAG}_bench.json
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/${TAG}_bench_reference.json 2>> $OUT/${TAG}_bench.err; echo "ref rc=$?"
timeout 300 python profiles/prof_step.py > $OUT/${TAG}_prof_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -s 42 -c 26 -f -o $OUT/${TAG}_full python profiles/prof_step.py > $OUT/${TAG}_ncu_full.log 2>&1
echo "ncu full rc=$?"; cat $OUT/${TAG}_prof_plain.log
if [ "${3:-}" == "no_launchlist" ]; then kill $SMI 2>/dev/null; ls -la $OUT | head -40; exit 0; fi
timeout 300 python bench.py --steps 2 --warmup 3 --no-configs --no-cpu-baseline > $OUT/${TAG}_bench_short.json 2>> $OUT/${TAG}_bench.err && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 100 -c 400 --csv --log-file $OUT/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-configs --no-cpu-baseline > $OUT/${TAG}_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
kill $SMI 2>/dev/null
ls -la $OUT | head -30
