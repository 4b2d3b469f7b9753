#!/bin/bash
# One bounded GPU session (edit per shot): parity tests of what changed, A/B of a tunable through bench.py.  Logs -> gpurun_out/.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_shard.py -m gpu -q -x > gpurun_out/t_parity.log 2>&1; echo "parity rc=$?"
tail -3 gpurun_out/t_parity.log
for it in 16 8 32; do
  FS3_SCAN_ITEMS=$it timeout 100 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_scan$it.json 2> gpurun_out/bench_scan$it.err; echo "bench scan=$it rc=$?"
done
