#!/bin/bash
# one ncu --set full capture of the two-step sweep (variant $1)
mkdir -p gpurun_out
V=${1:-10}
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fused2 -s 3 -c 1 -o gpurun_out/r01h_prof_fused2_v$V -f python bench.py --two-step --variant2 $V --steps 11 --warmup 3 --e2e-steps 0 --no-cpu-baseline > gpurun_out/r01h_ncu_full_v$V.log 2>&1; echo "ncu rc=$?"
