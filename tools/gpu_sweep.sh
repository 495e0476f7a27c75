#!/bin/bash
# tests of the fused kernels + variant sweep at 16384^2
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "fused or tmask or sw_steps or config3 or fullsize" > gpurun_out/pytest_fused.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_fused.log
tail -4 gpurun_out/pytest_fused.log
bash tools/sweep.sh "${1:-1 2 5 22 23 24 25 26 27 28 29 30}" 40 16384 | tee gpurun_out/sweep.txt
for pr in 0 1 2; do echo "L2PROMO=$pr"; DLESM_TMA_L2PROMO=$pr bash tools/sweep.sh "25" 40 16384; done | tee -a gpurun_out/sweep.txt
