#!/bin/bash
# ncu evidence for the bench (run via gpurun on one GPU). Plain run first, then the launch list, then one full capture.
mkdir -p gpurun_out
WL=${WL:-target}
CMD="python bench.py --steps 3 --warmup 3 --no-extras --workload $WL"
$CMD > gpurun_out/plain_$WL.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$WL.csv $CMD > gpurun_out/ncu_launches_$WL.log 2>&1
$CMD > gpurun_out/plain2_$WL.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_decode -s 3 -c 1 -o gpurun_out/prof_decode_$WL -f $CMD > gpurun_out/ncu_full_$WL.log 2>&1
tail -2 gpurun_out/ncu_full_$WL.log
