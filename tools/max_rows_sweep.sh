for MR in ${SWEEP:-18944 37888 56832 65536 18944}; do
  echo "== max_rows=$MR"
  python bench.py --max-rows $MR --steps 4 --warmup 3 --no-cpu-baseline --no-configs 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('value %.0f e2e %.0f ms %.1f frac %.4f clocks %s' % (d['value'], d['e2e']['value'], d['ms_per_step'], d['roofline']['frac'], d['clocks'].get('sm_mhz'))); print(d['breakdown_ms'])"
done
