#!/bin/bash
# config-5 microbench only (halo_exchange direct/NCCL + field_checksum) on N GPUs: tools/config5_only.sh N
n=$1; T=${TAG:-r02}
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29541 \
  bench.py --gpus $n --steps 4 --warmup 3 --no-parity --e2e-steps 0 --no-cpu-baseline --sustain-s 0 --no-strong --sweep one \
  > gpurun_out/${T}_config5_n$n.log 2>&1
echo "config5 rc=$?"
grep '^{' gpurun_out/${T}_config5_n$n.log | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
for c in d['config5']: print({k:(round(v,1) if isinstance(v,float) else v) for k,v in c.items()})" | tee gpurun_out/${T}_config5_n$n.txt
