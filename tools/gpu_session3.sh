#!/bin/bash
tag=${1:-r02p}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_shard.py tests/test_levelset_lines.py tests/test_state_file.py -m gpu -x -q > gpurun_out/${tag}_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_tests.log
tail -25 gpurun_out/${tag}_tests.log
