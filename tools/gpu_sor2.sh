#!/bin/bash
# sharded SOR on 2 GPUs: multi-process parity tests, then small slabs (128^3 / 2 exposes the per-half-sweep protocol cost) and 256^3
# usage: gpurun --gpus 2 -- bash tools/gpu_sor2.sh <tag> [grids...]
tag=${1:-sor2}; shift; GRIDS=${@:-128 192 256}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_shard.py -m gpu -x -q -k multiprocess > gpurun_out/${tag}_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_tests.log
tail -5 gpurun_out/${tag}_tests.log
for G in $GRIDS; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --grid $G --steps 10 --warmup 3 --no-cpu-baseline 2>gpurun_out/${tag}_g${G}.err | tee gpurun_out/${tag}_g${G}_n2.json | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('grid $G', round(d['ms_per_step'],3), d['check']['status'], 'sor_us', round(d['roofline']['avg_launch_us'],2), {k:round(v,2) for k,v in d['stage_ms_per_step'].items()})" || tail -5 gpurun_out/${tag}_g${G}.err
done
true
