#!/bin/bash
# sequences-per-warp sweep (SGB_GROUP) on small and large batches
for wl in ${WLS:-cfg2 target}; do
for g in ${GROUPS_:-32 16 8 4}; do
  out=$(SGB_GROUP=$g timeout 300 python bench.py --steps 10 --no-extras --workload $wl 2>&1 | tail -1)
  echo "$wl group=$g $(echo "$out" | python -c "import json,sys
try:
    d=json.loads(sys.stdin.read()); r=d['roofline']; print('kernel_ms %.4f pack_ms %.4f frac %.4f value %.4g e2e %.4g' % (r['kernel_ms'], r['pack_ms'], r['frac'], d['value'], d['e2e']['value']))
except Exception as e: print('FAILED', e)")"
done; done
