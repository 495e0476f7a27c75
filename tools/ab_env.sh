#!/bin/bash
# A/B of one environment knob on N GPUs: tools/ab_env.sh N VAR "v1 v2 ..." [bench args]
n=$1; var=$2; vals=$3; shift 3
T=${TAG:-r02}
mkdir -p gpurun_out
for v in $vals; do
  if [ $n -gt 1 ]; then
    env $var=$v timeout ${TMO:-400} python -m torch.distributed.run --nnodes=1 --nproc-per-node $n \
      --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $n --steps 100 --warmup 5 --no-parity --no-config5 \
      --e2e-steps 0 --no-cpu-baseline --sustain-s 0 "$@" > gpurun_out/${T}_ab_${var}_${v}_n$n.log 2>&1
  else
    env $var=$v timeout ${TMO:-400} python bench.py --steps 100 --warmup 5 --no-parity --no-config5 \
      --e2e-steps 0 --no-cpu-baseline --sustain-s 0 "$@" > gpurun_out/${T}_ab_${var}_${v}_n$n.log 2>&1
  fi
  echo "ab $var=$v rc=$?"
  grep '^{' gpurun_out/${T}_ab_${var}_${v}_n$n.log > gpurun_out/${T}_ab_${var}_${v}_n$n.json
done
