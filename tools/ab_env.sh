#!/bin/sh
# A / B / A of one environment switch on the headline step (same box, back to back): sh tools/ab_env.sh FB200_GEMM_RESPF 0 1
VAR=$1; A=$2; B=$3
for v in $A $B $A; do
  echo "== $VAR=$v"
  env $VAR=$v python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-configs 2>/dev/null | python -c "
import json,sys
for ln in sys.stdin:
    ln=ln.strip()
    if ln.startswith('{'):
        d=json.loads(ln); print('value %.0f e2e %.0f ms %.1f frac %.4f clocks %s parity %s' % (d['value'], d['e2e']['value'], d['ms_per_step'], d['roofline']['frac'], d['clocks'].get('sm_mhz'), d.get('parity',{}).get('ids_equal')))
"
done
