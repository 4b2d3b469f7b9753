set -x
run() { N=$1; G=$2; S=$3; W=$4; tag=$5
  if [ $N -eq 1 ]; then timeout 400 python bench.py --gpus 1 --grid $G --steps $S --warmup $W --no-cpu-baseline > gpurun_out/r01o_${tag}.json 2> gpurun_out/r01o_${tag}.err
  else timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --grid $G --steps $S --warmup $W --no-cpu-baseline > gpurun_out/r01o_${tag}.json 2> gpurun_out/r01o_${tag}.err; fi
  echo "$tag rc=$?"; tail -c 300 gpurun_out/r01o_${tag}.err | tail -3; }
run 8 512 3 2 g512_n8
run 1 512 3 2 g512_n1
run 1 256 5 3 g256_n1
nvidia-smi --query-gpu=memory.used --format=csv | head -3
