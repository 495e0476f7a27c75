#!/bin/bash
# usage: tools/gpu_scale.sh N [TAG]  -- bench at N GPUs: weak (16384^2 per GPU) and strong (16384^2 in total)
N=${1:-2}; TAG=${2:-r01g}
mkdir -p gpurun_out
run() { # $1 = scaling, $2 = extra args
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 \
      bench.py --gpus $N --steps 100 --warmup 5 --scaling $1 $2 > gpurun_out/${TAG}_bench_n${N}_$1.log 2>&1; echo "bench $1 rc=$?"
  grep '^{' gpurun_out/${TAG}_bench_n${N}_$1.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['scaling'], d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['halo_traffic'])"
}
run weak "--e2e-steps 2"
run strong "--e2e-steps 2"
