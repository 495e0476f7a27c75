#!/bin/bash
# A/B of the push hand-shake on N GPUs: in-sweep (default) vs one-warp kernel after every sweep.
#   tools/ab_push_sync.sh N [bench args]
n=${1:-4}; shift || true
T=${TAG:-r02}
mkdir -p gpurun_out
for mode in sweep kernel; do
  DLESM_PUSH_SYNC=$mode timeout ${TMO:-400} python -m torch.distributed.run --nnodes=1 --nproc-per-node $n \
    --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $n --steps 100 --warmup 5 --no-parity --no-config5 \
    --e2e-steps 0 --no-cpu-baseline --sustain-s 0 "$@" > gpurun_out/${T}_ab_${mode}_n$n.log 2>&1
  echo "ab $mode rc=$?"
  grep '^{' gpurun_out/${T}_ab_${mode}_n$n.log > gpurun_out/${T}_ab_${mode}_n$n.json
done
