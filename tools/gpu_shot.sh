#!/bin/bash
# One bounded GPU session: sharded (lockstep) parity tests after the sweep grid-size change.  Logs -> gpurun_out/.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 60 python -m pytest tests/test_gpu_shard.py tests/test_state_file.py -m gpu -q -x > gpurun_out/t_shard.log 2>&1; echo "shard tests rc=$?"; tail -2 gpurun_out/t_shard.log
