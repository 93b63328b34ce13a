#!/bin/bash
# Round 2, two-GPU call #2: data-parallel parity only (lr = 0 form of tools/dp_check.py).
mkdir -p gpurun_out
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29533 tools/dp_check.py > $O/r02g_dp_check.json 2> $O/r02g_dp_check.err; echo "dp_check rc=$?"; cat $O/r02g_dp_check.json; tail -3 $O/r02g_dp_check.err
