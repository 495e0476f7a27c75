#!/bin/bash
# full single-GPU pass: every gpu test, smoke, microbench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_full.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_full.log
tail -3 gpurun_out/pytest_gpu_full.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 600 python tools/microbench.py --sizes 4096,16384 > gpurun_out/micro_psy.jsonl 2> gpurun_out/micro_psy.err; echo "micro rc=$?"; tail -2 gpurun_out/micro_psy.err
python - <<PY
import json
for l in open("gpurun_out/micro_psy.jsonl"):
    d=json.loads(l); print(d["op"], d["local_tile"], round(d["us_per_call"],1), round(d.get("hbm_frac",0),3))
PY
