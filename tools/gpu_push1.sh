#!/bin/bash
# single GPU: halo push with self neighbours (parity) + the bench in both halo modes
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -q -x -k "push or step_host" > gpurun_out/pytest_push.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_push.log
tail -4 gpurun_out/pytest_push.log
for h in exchange push; do
  timeout 600 python bench.py --steps 100 --warmup 5 --no-cpu-baseline --e2e-steps 0 --halo $h > gpurun_out/bench_n1_$h.log 2>&1
  echo "halo $h rc=$?"; grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_n1_$h.log | head -1; grep -o '"frac": [0-9.]*' gpurun_out/bench_n1_$h.log
done
