#!/bin/bash
# usage: tools/gpu_multi.sh N   -- multi-rank parity worker + bench at N GPUs (push and exchange halos)
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus.txt
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 tests/dist_worker.py > gpurun_out/dist_worker_n$N.log 2>&1; echo "dist_worker rc=$?" | tee -a gpurun_out/dist_worker_n$N.log
grep -v "^\[W\|^W1\|OMP_NUM\|^\*\*\*" gpurun_out/dist_worker_n$N.log | tail -12
for h in exchange push; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $N --steps 100 --warmup 5 --halo $h > gpurun_out/bench_n${N}_$h.log 2>&1; echo "bench $h rc=$?"
  grep '^{' gpurun_out/bench_n${N}_$h.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['e2e']['value'])"
done
