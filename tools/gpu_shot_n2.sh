# 2-GPU session (gpurun --gpus 2 -- 'bash tools/gpu_shot_n2.sh'): the real 2-process parity test, then bench.py under torchrun.
# Set FS3_LS_BLOCKS=<n> to A/B the grid size of the replicated level-set sweeps.  Logs -> gpurun_out/.
mkdir -p gpurun_out
timeout 150 python -m pytest tests/test_gpu_shard.py::test_multiprocess_two_gpus -m gpu -q > gpurun_out/t_mp2.log 2>&1; echo "mp2 rc=$?"; tail -2 gpurun_out/t_mp2.log
NS=2 timeout 200 bash tools/scale_probe.sh 256 5 3 n2
