#!/bin/bash
# Round 2, third call: GPU suite incl. the distillation tests and the re-based sampler parity bars; per-pixel kernels.
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > $O/r02c_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -25 $O/r02c_pytest_gpu.log
timeout 300 python tools/profile_elementwise.py > $O/r02c_elementwise.json 2> $O/r02c_elementwise.err; echo "elementwise rc=$?"; cat $O/r02c_elementwise.json
