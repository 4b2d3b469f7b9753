#!/bin/bash
# One bounded GPU session: new-feature parity tests first, then the bench, the consumer microbench, the rest of the suite.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 240 python -m pytest tests/test_state_file.py tests/test_consumer.py -m gpu -q --deselect tests/test_consumer.py::test_gpu_consumer_at_256_properties > gpurun_out/t_new.log 2>&1; echo "new tests rc=$?"
tail -3 gpurun_out/t_new.log
timeout 150 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 90 python tools/consumer_bench.py --state /tmp/s256.fs3 > gpurun_out/consumer_bench.jsonl 2> gpurun_out/consumer_bench.err; echo "consumer bench rc=$?"
timeout 120 python -m pytest tests/test_consumer.py::test_gpu_consumer_at_256_properties -m gpu -q > gpurun_out/t_256.log 2>&1; echo "256 test rc=$?"
tail -3 gpurun_out/t_256.log
timeout 300 python -m pytest tests -m gpu -q -x --deselect tests/test_consumer.py::test_gpu_consumer_at_256_properties --ignore tests/test_state_file.py > gpurun_out/t_all.log 2>&1; echo "suite rc=$?"
tail -3 gpurun_out/t_all.log
timeout 100 ncu --set full --clock-control none -k regex:k_blur_phi_v4 -c 1 --csv --page raw --log-file gpurun_out/blur_ncu_raw.csv python tools/consumer_bench.py --reps 1 > gpurun_out/blur_ncu.log 2>&1; echo "ncu rc=$?"
