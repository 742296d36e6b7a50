# resize strip-height sweep (run under gpurun)
for t in ${@:-8 12 16 24 32 48}; do
  ORBX_RS_TARGET=$t python bench.py --steps 3 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('RS_TARGET', $t, 'value', round(d['value']), 'resize', round(d['stages']['resize']['ms_per_step'],3))"
done
