#!/bin/bash
# multi-GPU session: gpurun --gpus N -- bash tools/gpu_session_n.sh <tag> <N> [grids...]
tag=${1:-r02q}; N=${2:-2}; shift; shift; GRIDS=${@:-256}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/${tag}_gpus.txt
timeout 900 python -m pytest tests/test_gpu_shard.py -m gpu -x -q -k multiprocess > gpurun_out/${tag}_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_tests.log
tail -4 gpurun_out/${tag}_tests.log
for G in $GRIDS; do
steps=10; [ $G -ge 512 ] && steps=4
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --grid $G --steps $steps --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_g${G}_n${N}.json 2> gpurun_out/${tag}_g${G}_n${N}.err; echo "bench N=$N grid=$G rc=$?"
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${tag}_g${G}_n${N}.json").read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ("ms_per_step","value")}, d.get("check",{}).get("status"), "e2e", round(d["e2e"]["ms_per_step"],2), {k:round(v,3) for k,v in d["stage_ms_per_step"].items()}, "sor_us", round(d["roofline"]["avg_launch_us"],2))
except Exception as e:
    print("no json", e)
PY
tail -3 gpurun_out/${tag}_g${G}_n${N}.err | grep -v "OMP\|\*\*\*"
done
true
