#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "fused_variants_match or config3" > gpurun_out/pytest_fused.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_fused.log
tail -3 gpurun_out/pytest_fused.log
bash tools/sweep.sh "25 28 31 32 33 34" 40 16384 | tee gpurun_out/sweep.txt
for sr in 128 512 1024; do echo "SEG_ROWS=$sr"; DLESM_SEG_ROWS=$sr bash tools/sweep.sh "25 28 32" 40 16384; done | tee -a gpurun_out/sweep.txt
