#!/bin/bash
# sharded SOR on 4 GPUs: multi-process parity tests (world 2 and 4, fp64 and fp32), bench at 256^3 on 4 ranks and 128^3 on 2 (protocol floor)
# usage: gpurun --gpus 4 -- bash tools/gpu_sor4.sh <tag>     (time stamps: torchrun --nproc-per-node 4 tools/sor_trace.py 256)
tag=${1:-sor4}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_shard.py -m gpu -x -q -k "multiprocess and (2-24-4 or 4-12 or 4-64 or fp32)" 2>&1 | tail -2
for cfg in "4 256" "2 128"; do set -- $cfg
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $1 --grid $2 --steps 10 --warmup 3 --no-cpu-baseline 2>gpurun_out/${tag}_g$2_n$1.err | tee gpurun_out/${tag}_g$2_n$1.json | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('grid $2 N=$1', round(d['ms_per_step'],3), d['check']['status'], 'sor_us', round(d['roofline']['avg_launch_us'],2), {k:round(v,2) for k,v in d['stage_ms_per_step'].items()})" || tail -5 gpurun_out/${tag}_g$2_n$1.err
done
true
