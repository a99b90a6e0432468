#!/bin/bash
# Multi-GPU bench lines (run under `gpurun --gpus N`): weak-scaling headline, strong-scaling BASELINE configs[2] (cfg3) and [3] (cfg4).
#   tools/gpu_scale.sh N [workloads...]     e.g. tools/gpu_scale.sh 8 target cfg3:strong cfg4
N=$1; shift
mkdir -p gpurun_out
port=29600
for w in "$@"; do
  name=${w%%:*}; sc=""; tag=$name
  if [[ "$w" == *":strong" ]]; then sc="--scaling strong"; tag="${name}_strong"; fi
  port=$((port+1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port bench.py --gpus $N \
      --steps 20 --warmup 3 --workload $name $sc > gpurun_out/r02_scale_${tag}_${N}gpu.json 2> gpurun_out/r02_scale_${tag}_${N}gpu.err
  echo "$tag N=$N rc=$?"
  grep '^{' gpurun_out/r02_scale_${tag}_${N}gpu.json | python -c "
import sys, json
for ln in sys.stdin:
    d = json.loads(ln); print(d['config']['cells_per_gpu'], 'cells/gpu', d['scaling'], 'value', '%.4g' % d['value'], 'ms', '%.4f' % d['ms_per_step'], d['schedule'], d['gather'], 'e2e', '%.4g' % d['e2e']['value'], 'e2e_ms', '%.3f' % d['e2e']['ms_per_step'], d.get('multi_device_check', {}).get('equal_to_single_device'))"
done
