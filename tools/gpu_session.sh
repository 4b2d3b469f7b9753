#!/bin/bash
# One bounded single-GPU session: GPU test suite, bench, live SOR traffic.  Usage: gpurun --timeout 1500 -- bash tools/gpu_session.sh <tag>
tag=${1:-r02}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/${tag}_gpu.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q --durations=12 > gpurun_out/${tag}_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_tests.log
tail -25 gpurun_out/${tag}_tests.log
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/${tag}_bench.json
timeout 200 python tools/sor_traffic_gpm.py --out gpurun_out/${tag}_sor_traffic.json > gpurun_out/${tag}_gpm.log 2>&1; echo "gpm rc=$?"; tail -c 1500 gpurun_out/${tag}_gpm.log
timeout 100 python bench.py --impl reference --grid 96 --steps 2 --warmup 0 > gpurun_out/${tag}_ref.json 2>&1; tail -c 800 gpurun_out/${tag}_ref.json
