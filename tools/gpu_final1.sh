#!/bin/bash
# final single-GPU pass: every gpu test, smoke, default bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r01r_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r01r_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r01r_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r01r_smoke.log
timeout 900 python bench.py > gpurun_out/r01r_bench_n1.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/r01r_bench_n1.log | cut -c1-260
