#!/bin/bash
tag=${1:-r02t}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_tests.log
tail -5 gpurun_out/${tag}_tests.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/${tag}_bench.json").read().strip().splitlines()[-1])
print({k:d.get(k) for k in ("ms_per_step","value")}, "e2e", d["e2e"]["ms_per_step"], {k:round(v,3) for k,v in d["stage_ms_per_step"].items()}, "sor_us", round(d["roofline"]["avg_launch_us"],2))
PY
