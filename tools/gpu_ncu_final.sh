#!/bin/bash
# ncu --set full of the one-step sweep as shipped, and of the masked two-step sweep (config 3 pattern)
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fused_tma_kernel -s 4 -c 1 -o gpurun_out/r01z_prof_fused1 -f python bench.py --sweep one --steps 9 --warmup 3 --e2e-steps 0 --no-cpu-baseline > gpurun_out/r01z_ncu1.log 2>&1; echo "ncu one rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fused2 -s 1 -c 1 -o gpurun_out/r01z_prof_fused2s_masked -f python bench.py --sweep two --tmask --steps 9 --warmup 3 --e2e-steps 0 --no-cpu-baseline > gpurun_out/r01z_ncu2.log 2>&1; echo "ncu masked rc=$?"
