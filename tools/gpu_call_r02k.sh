#!/bin/bash
# GroupNorm CTA-shape sweep after the round-2 kernel changes (slice width x threads)
mkdir -p gpurun_out
O=gpurun_out
for T in 512 256; do for CS in 16 32; do
  echo "== B200_GN_THREADS=$T B200_GN_CS=$CS"
  B200_GN_THREADS=$T B200_GN_CS=$CS timeout 200 python tools/time_gn.py 2>&1 | grep -E "HW784|HW196"
done; done > $O/r02k_gn_sweep.txt 2>&1
cat $O/r02k_gn_sweep.txt
