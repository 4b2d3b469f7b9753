# 2-GPU session: level-set sweep grid size A/B on the sharded step (bench.py under torchrun).  Logs -> gpurun_out/.
mkdir -p gpurun_out
for B in 222 185 148; do
  FS3_LS_BLOCKS=$B NS=2 timeout 100 bash tools/scale_probe.sh 256 5 3 r01af_lsb$B
done
