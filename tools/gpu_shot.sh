#!/bin/bash
# One bounded GPU session (edit per shot): parity tests of what changed, microbench, ncu captures.  Logs -> gpurun_out/.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 200 python -m pytest tests/test_consumer.py -m gpu -q > gpurun_out/t_new.log 2>&1; echo "consumer tests rc=$?"
tail -3 gpurun_out/t_new.log
timeout 90 python tools/consumer_bench.py > gpurun_out/consumer_bench.jsonl 2> gpurun_out/consumer_bench.err; echo "consumer bench rc=$?"
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 520 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; echo "ncu launch list rc=$?"
timeout 100 ncu --set full --clock-control none -k regex:k_blur_phi_v4 -c 1 --csv --page raw --log-file gpurun_out/blur_ncu_raw.csv python tools/consumer_bench.py --reps 1 > gpurun_out/blur_ncu.log 2>&1; echo "ncu blur rc=$?"
