#!/bin/bash
# first GPU pass of the round: probes, gpu tests, bench, ncu launch list, ncu full capture
mkdir -p gpurun_out
{ nproc; free -g | head -2; gfortran --version 2>&1 | head -1; mpirun --version 2>&1 | head -1; nvidia-smi -L; } > gpurun_out/probe.txt 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_default.log 2>&1; echo "bench rc=$?" >> gpurun_out/bench_default.log
python bench.py --impl reference --steps 6 --warmup 1 > gpurun_out/bench_ref.log 2>&1
CMD="python bench.py --steps 5 --warmup 3 --e2e-steps 1 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fused_tma -s 4 -c 2 -o gpurun_out/prof_fused -f $CMD > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/bench_default.log; tail -1 gpurun_out/bench_ref.log
