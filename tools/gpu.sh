#!/bin/bash
# One parameterised GPU-box runner (replaces the one-off gpu_*.sh scripts of round 1).
#   tools/gpu.sh tests            # every -m gpu test + smoke
#   tools/gpu.sh bench [args..]   # python bench.py args..       (1 GPU)
#   tools/gpu.sh benchn N [args]  # torchrun bench.py --gpus N args..
#   tools/gpu.sh worker N         # tests/dist_worker.py on N GPUs
#   tools/gpu.sh launches [args]  # ncu launch list of `bench.py args` (after a plain run exited 0)
#   tools/gpu.sh ncu REGEX TAG [args]  # ncu --set full of kernels matching REGEX in `bench.py args`
#   tools/gpu.sh micro [args]     # tools/microbench.py
# Outputs land in gpurun_out/ (merged back by gpurun).  TAG env prefixes file names.
set -u
mkdir -p gpurun_out
T=${TAG:-r02}
what=${1:-tests}; shift || true
trun() { n=$1; shift; timeout ${TMO:-900} python -m torch.distributed.run --nnodes=1 --nproc-per-node $n \
         --master-addr 127.0.0.1 --master-port ${PORT:-29533} "$@"; }
case $what in
  tests)
    timeout ${TMO:-1500} python -m pytest tests -m gpu -q "$@" > gpurun_out/${T}_pytest_gpu.log 2>&1
    echo "pytest rc=$?" | tee -a gpurun_out/${T}_pytest_gpu.log; tail -4 gpurun_out/${T}_pytest_gpu.log
    timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1
    echo "smoke rc=$?"; tail -1 gpurun_out/${T}_smoke.log ;;
  bench)
    timeout ${TMO:-900} python bench.py "$@" > gpurun_out/${T}_bench.log 2>&1; echo "bench rc=$?"
    grep '^{' gpurun_out/${T}_bench.log > gpurun_out/${T}_bench.json; tail -c 1500 gpurun_out/${T}_bench.log ;;
  benchn)
    n=$1; shift
    trun $n bench.py --gpus $n "$@" > gpurun_out/${T}_bench_n$n.log 2>&1; echo "bench n=$n rc=$?"
    grep '^{' gpurun_out/${T}_bench_n$n.log > gpurun_out/${T}_bench_n$n.json; tail -c 1500 gpurun_out/${T}_bench_n$n.log ;;
  worker)
    n=$1; shift
    trun $n tests/dist_worker.py "$@" > gpurun_out/${T}_dist_worker_n$n.log 2>&1
    echo "dist_worker rc=$?" | tee -a gpurun_out/${T}_dist_worker_n$n.log
    grep -v "^\[W\|^W1\|OMP_NUM\|^\*\*\*" gpurun_out/${T}_dist_worker_n$n.log | tail -8 ;;
  launches)
    CMD="python bench.py --steps 6 --warmup 3 --e2e-steps 0 --no-cpu-baseline --no-parity --no-config5 --no-strong $*"
    $CMD > gpurun_out/${T}_plain.log 2>&1 &&
    ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
        --log-file gpurun_out/${T}_launches.csv $CMD > gpurun_out/${T}_ncu_launches.log 2>&1
    echo "launch list rc=$?" ;;
  ncu)
    re=$1; tag=$2; shift 2
    CMD="python bench.py --steps 8 --warmup 3 --e2e-steps 0 --no-cpu-baseline --no-parity --no-config5 --no-strong $*"
    $CMD > gpurun_out/${T}_${tag}_plain.log 2>&1 &&
    timeout ${TMO:-900} ncu --set full --clock-control none --import-source on -k regex:$re -s ${SKIP:-2} -c 1 \
        -o gpurun_out/${T}_${tag} -f $CMD > gpurun_out/${T}_${tag}_ncu.log 2>&1
    echo "ncu $tag rc=$?" ;;
  nvlink)
    # NVLink counters of the push sweep on rank 0 of an N-rank run (N = $1)
    n=$1; shift
    ncu --query-metrics 2>/dev/null | grep -i "nvl" > gpurun_out/${T}_nvl_metrics.txt
    export TAG=$T NCU_METRICS=${NCU_METRICS:-nvltx__bytes.sum,nvlrx__bytes.sum,gpu__time_duration.sum}
    TMO=${TMO:-240} timeout ${TMO:-240} python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 \
        --master-port ${PORT:-29544} --no-python tools/ncu_rank0.sh bench.py --gpus $n --steps 8 --warmup 3 --e2e-steps 0 \
        --no-cpu-baseline --no-parity --no-config5 --no-strong --no-config3 --sustain-s 0 "$@" > gpurun_out/${T}_nvlink.log 2>&1
    echo "nvlink rc=$?"; wc -l gpurun_out/${T}_nvl_metrics.txt ;;
  probe)
    { nvidia-smi topo -m; echo; cat /sys/devices/system/node/online; grep -H . /sys/devices/system/node/node*/cpulist
      for d in /sys/bus/pci/devices/*; do c=$(cat $d/class 2>/dev/null); case $c in 0x0302*|0x0300*) echo "$d numa=$(cat $d/numa_node) $(cat $d/vendor)";; esac; done
      cat /sys/fs/cgroup/cpuset.cpus.effective /sys/fs/cgroup/cpuset.mems.effective 2>/dev/null; nproc; free -g | head -2
      grep -i "Mems_allowed_list\|Cpus_allowed_list" /proc/self/status; } > gpurun_out/${T}_probe.txt 2>&1
    echo "probe rc=$?" ;;
  micro)
    timeout ${TMO:-900} python tools/microbench.py "$@" > gpurun_out/${T}_micro.jsonl 2> gpurun_out/${T}_micro.err
    echo "micro rc=$?"; tail -2 gpurun_out/${T}_micro.err ;;
  *) echo "unknown action $what"; exit 2 ;;
esac
