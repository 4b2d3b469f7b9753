#!/bin/bash
tag=${1:-r02b}
mkdir -p gpurun_out
timeout 300 python tools/ls_bench.py 256 5 > gpurun_out/${tag}_lsbench.log 2>&1; tail -5 gpurun_out/${tag}_lsbench.log
timeout 100 python tools/ls_bench.py 128 5 >> gpurun_out/${tag}_lsbench.log 2>&1; tail -2 gpurun_out/${tag}_lsbench.log
timeout 900 python -m pytest tests/test_gpu_shard.py tests/test_levelset_lines.py tests/test_edge_scenes.py tests/test_state_file.py tests/test_cpp_host.py -m gpu -x -q > gpurun_out/${tag}_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_tests.log
tail -25 gpurun_out/${tag}_tests.log
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "not 256 and not 128" > gpurun_out/${tag}_tests2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_tests2.log
tail -15 gpurun_out/${tag}_tests2.log
