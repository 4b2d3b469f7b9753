#!/bin/bash
mkdir -p gpurun_out
for cfg in "24 4 4.0" "48 4 4.0" "24 6 2.0"; do set -- $cfg
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29655 tests/shard_worker.py $1 $2 --zflow $3 2>&1 | grep "rank 0\|SHARD_WORKER_OK\|Error" | head -5
done
