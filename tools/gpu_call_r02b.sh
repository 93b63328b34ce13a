#!/bin/bash
# Round 2, second call: re-run the GPU suite after the oracle-device / logdir fixes, the per-pixel kernels after their
# unrolling, and the per-layer conv timings of this repo next to cuDNN's (torch eager, channels_last, benchmark mode).
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s > $O/r02b_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -12 $O/r02b_pytest_gpu.log
timeout 300 python tools/profile_elementwise.py > $O/r02b_elementwise.json 2> $O/r02b_elementwise.err; echo "elementwise rc=$?"; cat $O/r02b_elementwise.json
timeout 600 python tools/sampler_parity.py --out $O/r02b_sampler_parity.json > $O/r02b_sampler_parity.txt 2>&1; echo "parity rc=$?"
timeout 300 python tools/time_cudnn_conv.py > $O/r02b_cudnn_conv.txt 2>&1; echo "cudnn rc=$?"
timeout 300 python tools/time_conv.py > $O/r02b_time_conv.txt 2>&1; echo "time_conv rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"q_sample|loss_kernel|sampler_step" -s 3 -c 3 -f -o $O/r02b_prof_elementwise python tools/profile_elementwise.py --reps 1 > $O/r02b_ncu_elementwise.log 2>&1; echo "ncu elementwise rc=$?"
