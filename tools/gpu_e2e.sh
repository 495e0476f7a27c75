#!/bin/bash
# banded host step: parity + e2e timing at a few band heights
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -q -x -k "step_host" > gpurun_out/pytest_step_host.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_step_host.log
tail -3 gpurun_out/pytest_step_host.log
for br in 0 64 256 1024; do
  python bench.py --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 3 --band-rows $br > gpurun_out/e2e_band$br.log 2>&1
  echo "band $br rc=$?"; grep -o '"e2e": {[^}]*}[^}]*}' gpurun_out/e2e_band$br.log | cut -c1-400
done
