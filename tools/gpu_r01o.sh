#!/bin/bash
# round-end evidence on one GPU: default bench, reference arm, ncu launch list of the same command shape,
# ncu --set full of the two-step sweep (the dominant kernel of the default run)
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/r01o_bench_n1.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/r01o_bench_n1.log | cut -c1-300
timeout 600 python bench.py --impl reference --steps 6 --warmup 1 > gpurun_out/r01o_bench_ref.log 2>&1; echo "ref rc=$?"
CMD="python bench.py --steps 6 --warmup 3 --e2e-steps 1 --no-cpu-baseline"
$CMD > gpurun_out/r01o_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01o_launches_bench_steps6.csv $CMD > gpurun_out/r01o_ncu_launches.log 2>&1; echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fused2 -s 1 -c 1 -o gpurun_out/r01o_prof_fused2s -f python bench.py --sweep two --steps 9 --warmup 3 --e2e-steps 0 --no-cpu-baseline > gpurun_out/r01o_ncu_full.log 2>&1; echo "ncu full rc=$?"
