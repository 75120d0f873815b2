for i in 1 2 3; do
for m in 0 1; do
  if [ $m = 1 ]; then export VXORB_NO_FLAT_UPLOAD=1; else unset VXORB_NO_FLAT_UPLOAD; fi
  python bench.py --no-cpu-baseline 2>/dev/null | python -c "
import json,sys,os
d=json.loads(sys.stdin.read()); print('strided  ' if os.environ.get('VXORB_NO_FLAT_UPLOAD') else 'flat     ', round(d['value']), round(d['e2e']['value']), round(d['e2e']['h2d_gbs'],1))"
done; done
