#!/bin/bash
# multi-GPU session: gpurun --gpus N -- bash tools/gpu_session_n.sh <tag> <N> [grid]
tag=${1:-r02q}; N=${2:-2}; G=${3:-256}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/${tag}_gpus.txt
timeout 900 python -m pytest tests/test_gpu_shard.py -m gpu -x -q -k multiprocess > gpurun_out/${tag}_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_tests.log
tail -6 gpurun_out/${tag}_tests.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --grid $G --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_bench_n${N}.json 2> gpurun_out/${tag}_bench_n${N}.err; echo "bench rc=$?"
tail -c 2500 gpurun_out/${tag}_bench_n${N}.json; tail -5 gpurun_out/${tag}_bench_n${N}.err
