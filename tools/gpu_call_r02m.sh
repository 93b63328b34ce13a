#!/bin/bash
# The reference's default batch (gms.main --bs 64) on one B200: our step, the reference in torch eager, and where the
# time goes per launch.
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python bench.py --batch 64 --steps 50 --warmup 10 --sample-n 64 --sample-steps 200 --gpu-baseline 0 --small-n 0 --cpu-steps 1 > $O/r02m_bench_b64.json 2> $O/r02m_bench_b64.err; echo "bench b64 rc=$?"; tail -2 $O/r02m_bench_b64.err
timeout 600 python tools/gpu_torch_baseline.py --batch 64 --steps 20 --warmup 5 --sample-n 64 --sample-steps 20 --variants tf32,bf16,tf32_cl,bf16_cl > $O/r02m_gpu_torch_baseline_b64.json 2> $O/r02m_gpu_torch_baseline_b64.err; echo "torch b64 rc=$?"; tail -2 $O/r02m_gpu_torch_baseline_b64.err
timeout 300 python tools/per_op.py --batch 64 --csv $O/r02m_per_op_b64.csv > $O/r02m_per_op_b64.txt 2>&1; echo "per_op rc=$?"; tail -5 $O/r02m_per_op_b64.txt
