# bench.py at N = 1, 2, 4, 8 on one box (256^3) and, with "512" as first argument, 512^3 at N = 8
G=${1:-256}; S=${2:-5}; W=${3:-3}; TAG=${4:-probe}
for N in ${NS:-1 2 4 8}; do
  if [ $N -eq 1 ]; then timeout 400 python bench.py --gpus 1 --grid $G --steps $S --warmup $W --no-cpu-baseline > gpurun_out/${TAG}_g${G}_n$N.json 2> gpurun_out/${TAG}_g${G}_n$N.err
  else timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --grid $G --steps $S --warmup $W --no-cpu-baseline > gpurun_out/${TAG}_g${G}_n$N.json 2> gpurun_out/${TAG}_g${G}_n$N.err; fi
  echo "grid $G N=$N rc=$?"
  python - <<P
import json
try:
    for l in open('gpurun_out/${TAG}_g${G}_n$N.json'):
        if l.startswith('{'):
            d=json.loads(l); print(round(d['ms_per_step'],2), 'e2e', round(d['e2e']['ms_per_step'],1), 'sor us', round(d['roofline']['avg_launch_us'],1), {k:round(v,2) for k,v in d['stage_ms_per_step'].items()})
except Exception as e: print('parse fail', e)
P
done
