#!/bin/bash
# One bounded GPU session: edge-scene parity tests.  Logs -> gpurun_out/.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 24 python -m pytest tests/test_edge_scenes.py -m gpu -q > gpurun_out/t_edge.log 2>&1; echo "edge tests rc=$?"; tail -15 gpurun_out/t_edge.log
