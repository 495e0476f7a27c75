#!/bin/bash
# tools/run_trace.sh N  -- CTA timelines of the strong-scaling tiles on N GPUs (push vs exchange, two-step / one-step)
n=${1:-4}; T=${TAG:-r02}
mkdir -p gpurun_out
out=gpurun_out/${T}_cta_trace_n$n.jsonl; : > $out
for sweep in two one; do for halo in push exchange; do
  if [ $n -gt 1 ]; then
    timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29547 \
      tools/cta_trace.py --size ${SIZE:-16384} --scaling ${SCALING:-strong} --sweep $sweep --halo $halo > gpurun_out/_tr.log 2>&1
  else
    timeout 200 python tools/cta_trace.py --size ${SIZE:-4096} --sweep $sweep --halo $halo > gpurun_out/_tr.log 2>&1
  fi
  echo "trace $sweep $halo rc=$?"
  grep '^{' gpurun_out/_tr.log >> $out || tail -5 gpurun_out/_tr.log
done; done
