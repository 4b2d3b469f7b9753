#!/bin/bash
tag=${1:-r02e}
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "stages_from_identical_state or golden or free_running" > gpurun_out/${tag}_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_tests.log; tail -5 gpurun_out/${tag}_tests.log
for e in 0; do timeout 200 python tools/ls_trace.py 256 8 $e 2>&1 | grep -v "nvcc warning" | head -16 >> gpurun_out/${tag}_lstrace.log; done
cat gpurun_out/${tag}_lstrace.log
timeout 100 python tools/ls_bench.py 256 3 2>&1 | tail -2
