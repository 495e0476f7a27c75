#!/bin/bash
# two-step push trace on N GPUs with the env given on the command line, tag in $1
tag=$1; n=$2; shift 2
timeout 200 env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29547 \
  tools/cta_trace.py --size ${SIZE:-16384} --scaling strong --sweep two --halo push > gpurun_out/_tr.log 2>&1
echo "trace $tag rc=$?"; grep '^{' gpurun_out/_tr.log > gpurun_out/${tag}.jsonl
