#!/bin/bash
# default bench (both sweeps), ncu launch list of the same command shape, full gpu tests
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/r01i_bench_n1.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/r01i_bench_n1.log | cut -c1-600
CMD="python bench.py --steps 6 --warmup 3 --e2e-steps 1 --no-cpu-baseline"
$CMD > gpurun_out/r01i_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01i_launches_bench_steps6.csv $CMD > gpurun_out/r01i_ncu_launches.log 2>&1; echo "launch list rc=$?"
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r01i_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r01i_pytest_gpu.log
