#!/bin/bash
tag=${1:-r02e}
mkdir -p gpurun_out
for bz in 8 16 4; do timeout 200 python tools/ls_trace.py 256 $bz 2>&1 | grep -v "nvcc warning" | head -6 >> gpurun_out/${tag}_lstrace.log; done
cat gpurun_out/${tag}_lstrace.log
