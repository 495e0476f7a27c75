#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --size 1024 --steps 20 --warmup 3 --e2e-steps 0 --no-cpu-baseline"
$CMD > gpurun_out/r01p_plain.log 2>&1
ncu --metrics gpu__time_duration.sum,launch__grid_size --clock-control none -c 200 --csv --log-file gpurun_out/r01p_launches_1024.csv $CMD > gpurun_out/r01p_ncu.log 2>&1; echo rc=$?
