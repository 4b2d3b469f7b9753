#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 9 python tests/gls_worker.py 16 24 > gpurun_out/gls_worker.log 2>&1; echo "gls worker rc=$?"; tail -4 gpurun_out/gls_worker.log
