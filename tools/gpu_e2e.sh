#!/bin/bash
mkdir -p gpurun_out
for br in 0 64 256 512 1024; do
  timeout 300 python bench.py --sweep one --steps 6 --warmup 3 --e2e-steps 4 --band-rows $br --no-cpu-baseline > gpurun_out/r01t_e2e.log 2>&1
  grep '^{' gpurun_out/r01t_e2e.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); e=d['e2e']; print('band $br', round(e['ms_per_step'],1), round(e['pcie_GBs_per_direction_achieved'],1), round(e['pcie_GBs_per_direction_both_busy_measured'],1), round(e['frac_of_pcie'],3))"
done
