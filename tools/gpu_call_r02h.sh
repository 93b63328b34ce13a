#!/bin/bash
# Round 2: attention extension + 3-channel images — kernel tests first, then the whole GPU suite.
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests/test_gpu_attention.py -x -q > $O/r02h_pytest_attn.log 2>&1; echo "attn rc=$?"; tail -30 $O/r02h_pytest_attn.log
timeout 1500 python -m pytest tests -m gpu -q > $O/r02h_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -12 $O/r02h_pytest_gpu.log
