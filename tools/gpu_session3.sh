#!/bin/bash
tag=${1:-r02t}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_edge_scenes.py tests/test_gpu_shard.py tests/test_state_file.py tests/test_levelset_lines.py -m gpu -x -q -k "not 256" > gpurun_out/${tag}_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_tests.log
tail -12 gpurun_out/${tag}_tests.log
timeout 200 python tools/ls_bench.py 256 2 2>&1 | head -2
