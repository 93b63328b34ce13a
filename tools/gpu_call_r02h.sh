#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests/test_gpu_attention.py tests/test_gpu_model.py -q -k "attention or three_channel" > $O/r02h_pytest_attn.log 2>&1; echo "attn rc=$?"; tail -30 $O/r02h_pytest_attn.log
timeout 600 python bench.py --config 4 --steps 5 --warmup 3 --sample-steps 20 --bf16-leg 0 > $O/r02h_bench_config4.json 2> $O/r02h_bench_config4.err; echo "config4 rc=$?"; tail -3 $O/r02h_bench_config4.err; head -c 1500 $O/r02h_bench_config4.json
