#!/bin/bash
# A/B experimental builds on one GPU box: for every libsignals_b200*.so and staging mode, a short bench.
mkdir -p gpurun_out
: > gpurun_out/variants.txt
for lib in signals_b200/libsignals_b200*.so; do
  for stg in ${STAGES:-raw tma ldg}; do
    out=$(SGB_LIB=$PWD/$lib SGB_STAGE=$stg timeout 300 python bench.py --steps ${STEPS:-10} --no-extras --workload ${WL:-target} 2>&1 | tail -1)
    echo "$lib $stg $(echo "$out" | python -c "import json,sys
try:
    d=json.loads(sys.stdin.read()); r=d['roofline']; print('kernel_ms %.4f pack_ms %.4f frac %.4f value %.4g e2e %.4g' % (r['kernel_ms'], r['pack_ms'], r['frac'], d['value'], d['e2e']['value']))
except Exception as e: print('FAILED', e)")" | tee -a gpurun_out/variants.txt
  done
done
