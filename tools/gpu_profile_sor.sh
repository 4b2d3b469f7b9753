#!/bin/bash
# Short single-GPU profiling session (B200_PROFILING.md recipe): launch list of a short bench run + ncu --set full of two SOR half sweeps.
tag=${1:-r04}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
$CMD > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 150 -c 800 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu1.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_sor_half -s 450 -c 2 -f -o gpurun_out/${tag}_sor $CMD > gpurun_out/${tag}_ncu3.log 2>&1; echo "sor rc=$?"
ls -la gpurun_out | grep ${tag}
