#!/bin/bash
# ring-depth sensitivity of the strip kernel (how much shared memory can the epilogue staging take from the weight ring?)
mkdir -p gpurun_out
O=gpurun_out
for R in 10 3 2; do
  echo "== B200_STRIP_RING=$R"
  B200_STRIP_RING=$R timeout 300 python tools/time_conv.py 2>&1 | grep -E "R=3 mode=[01] (res=[01]|bf16)" | grep -E "H=28|H=14"
done > $O/r02i_ring.txt 2>&1
cat $O/r02i_ring.txt
