#!/bin/bash
tag=${1:-r02l}
mkdir -p gpurun_out
python tools/ls_ncu.py 128 > gpurun_out/${tag}_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_ls_sweep_strips -s 1 -c 1 -f -o gpurun_out/${tag}_ls python tools/ls_ncu.py 128 > gpurun_out/${tag}_ncu.log 2>&1
tail -5 gpurun_out/${tag}_ncu.log; ls -la gpurun_out/${tag}_ls.ncu-rep
