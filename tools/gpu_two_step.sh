#!/bin/bash
# two-step sweep: parity tests, bench per variant / segment height, one ncu capture
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -q -k "two_step" > gpurun_out/r01h_pytest_two_step.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r01h_pytest_two_step.log
B="python bench.py --two-step --steps 101 --warmup 5 --e2e-steps 0 --no-cpu-baseline"
show() { grep '^{' $1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$2', round(d['value']/1e9,1), round(d['ms_per_step'],3), round(d['roofline']['kernel_ms'],3), round(d['roofline']['frac'],3), d['clocks']['sm_mhz'], d['clocks']['reasons'])" || tail -3 $1; }
for v in ${VARIANTS:-1 2 3 4 5 7}; do
  timeout 300 $B --variant2 $v > gpurun_out/r01h_bench_two_step_v$v.log 2>&1; show gpurun_out/r01h_bench_two_step_v$v.log v$v
done
for sr in ${SEGS:-28 124 252}; do
  DLESM_SEG_ROWS2=$sr timeout 300 $B --variant2 2 > gpurun_out/r01h_bench_two_step_seg$sr.log 2>&1; show gpurun_out/r01h_bench_two_step_seg$sr.log seg$sr
done
if [ -n "$NCU" ]; then
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fused2_tma -s 3 -c 1 -o gpurun_out/r01h_prof_fused2 -f python bench.py --two-step --variant2 ${NCUV:-2} --steps 11 --warmup 3 --e2e-steps 0 --no-cpu-baseline > gpurun_out/r01h_ncu_full.log 2>&1; echo "ncu rc=$?"
fi
