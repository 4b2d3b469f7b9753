#!/bin/bash
tag=${1:-r02c}
mkdir -p gpurun_out
timeout 300 python tools/ls_bench.py 256 5 > gpurun_out/${tag}_lsbench.log 2>&1; tail -3 gpurun_out/${tag}_lsbench.log
timeout 100 python tools/ls_bench.py 128 5 >> gpurun_out/${tag}_lsbench.log 2>&1; tail -2 gpurun_out/${tag}_lsbench.log
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_edge_scenes.py -m gpu -x -q -k "not 256" > gpurun_out/${tag}_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_tests.log
tail -8 gpurun_out/${tag}_tests.log
