#!/bin/bash
# ncu evidence for the bench (run via gpurun on one GPU). Plain run first, then the launch list, then one full capture.
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-extras --workload ${WL:-target}"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_decode -s 3 -c 1 -o gpurun_out/prof_decode -f $CMD > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log; ls -la gpurun_out/
