#!/bin/bash
# re-entry verification pass: every gpu test, smoke, default bench, reference arm
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r01g_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r01g_pytest_gpu.log
tail -3 gpurun_out/r01g_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r01g_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r01g_smoke.log
timeout 900 python bench.py > gpurun_out/r01g_bench_n1.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/r01g_bench_n1.log
timeout 600 python bench.py --impl reference --steps 6 --warmup 1 > gpurun_out/r01g_bench_ref.log 2>&1; tail -1 gpurun_out/r01g_bench_ref.log
