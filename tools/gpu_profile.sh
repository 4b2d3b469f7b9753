#!/bin/bash
# One single-GPU profiling session (B200_PROFILING.md recipe): launch list of a short bench run, ncu --set full of the top kernels.
tag=${1:-r02}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
$CMD > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 150 -c 800 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu1.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_p2g_units -s 3 -c 3 -f -o gpurun_out/${tag}_p2g $CMD > gpurun_out/${tag}_ncu2.log 2>&1; echo "p2g rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_sor_half -s 450 -c 2 -f -o gpurun_out/${tag}_sor $CMD > gpurun_out/${tag}_ncu3.log 2>&1; echo "sor rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"k_advect|k_g2p|k_bin_rank_gather|k_scan_lookback|k_ls_seed|k_ex_fill_layers" -s 12 -c 7 -f -o gpurun_out/${tag}_particles $CMD > gpurun_out/${tag}_ncu4.log 2>&1; echo "particles rc=$?"
timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --gpu-formulation > gpurun_out/${tag}_gpuform.json 2>&1; echo "gpu-formulation rc=$?"; tail -c 600 gpurun_out/${tag}_gpuform.json
ls -la gpurun_out | grep ${tag}
