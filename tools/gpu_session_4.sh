#!/bin/bash
# 4-GPU session: multi-process shard tests (world 2/4), bench at N = 4 (256^3 and 512^3) and N = 2 (256^3).
tag=${1:-r02z}
mkdir -p gpurun_out
nvidia-smi -L | wc -l > gpurun_out/${tag}_gpus.txt
timeout 900 python -m pytest tests/test_gpu_shard.py -m gpu -x -q -k multiprocess -v > gpurun_out/${tag}_mp_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_mp_tests.log
grep -E "PASSED|FAILED|SKIPPED|passed|failed|rc=" gpurun_out/${tag}_mp_tests.log | tail -8
for cfg in "4 256 10" "4 512 4"; do set -- $cfg
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $1 --grid $2 --steps $3 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_g$2_n$1.json 2> gpurun_out/${tag}_g$2_n$1.err; echo "bench N=$1 grid=$2 rc=$?"
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${tag}_g$2_n$1.json").read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ("ms_per_step","value")}, d.get("check",{}).get("status"), {k:round(v,3) for k,v in d["stage_ms_per_step"].items()}, "sor_us", round(d["roofline"]["avg_launch_us"],2))
except Exception as e:
    print("no json", e)
PY
tail -3 gpurun_out/${tag}_g$2_n$1.err | grep -v "OMP\|\*\*\*"
done
