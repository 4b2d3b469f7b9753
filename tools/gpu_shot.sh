#!/bin/bash
# One bounded GPU session: the whole GPU suite, smoke(), the default bench.  Logs -> gpurun_out/.
#   gpurun --timeout 400 -- 'bash tools/gpu_shot.sh'
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 300 python -m pytest tests -m gpu -q -x > gpurun_out/t_all.log 2>&1; echo "suite rc=$?"; tail -2 gpurun_out/t_all.log
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 150 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
