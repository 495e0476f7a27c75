#!/bin/bash
mkdir -p gpurun_out
for sr in 16 24 32 40 48; do echo "SEG_ROWS=$sr"; DLESM_SEG_ROWS=$sr bash tools/sweep.sh "22 24 25 26 27 29 30" 40 16384; done | tee gpurun_out/sweep.txt
