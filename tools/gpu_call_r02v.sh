#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests/test_gpu_dropout.py -m gpu -q > $O/r02v_dropout.log 2>&1; echo "dropout rc=$?"; tail -30 $O/r02v_dropout.log | cut -c1-220
timeout 900 python -m pytest tests -m gpu -x -q > $O/r02v_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/r02v_pytest.log
