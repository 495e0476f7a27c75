#!/bin/bash
# torchrun --no-python tools/ncu_rank0.sh <bench args>: rank 0 runs under ncu (few metrics, one pass, only the
# sweep kernels -- they never wait on a peer, so profiling them cannot dead-lock the hand-shake), the others plain.
if [ "${LOCAL_RANK:-0}" = "0" ]; then
  exec ncu --metrics ${NCU_METRICS} --clock-control none -k regex:${NCU_KERNEL:-fused} -s ${NCU_SKIP:-6} -c ${NCU_COUNT:-4} --csv \
       --log-file gpurun_out/${TAG:-r02}_nvlink_rank0.csv python "$@"
else
  exec python "$@"
fi
