#!/bin/bash
# Round 2, first gpurun call: full GPU suite, default bench, both comparison arms, the sampler parity table, and the ncu
# captures r01's verdict found missing (GroupNorm backward, the per-pixel diffusion kernels at n = 8192).
# (B200_PROFILING.md: every ncu run follows a plain run of the same command that exited 0.)
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s > $O/r02a_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 $O/r02a_pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 > $O/r02a_bench_n1.json 2> $O/r02a_bench_n1.err; echo "bench rc=$?"; tail -3 $O/r02a_bench_n1.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $O/r02a_bench_reference.json 2> $O/r02a_bench_reference.err; echo "ref rc=$?"
timeout 600 python tools/sampler_parity.py --out $O/r02a_sampler_parity.json > $O/r02a_sampler_parity.txt 2>&1; echo "parity rc=$?"
timeout 900 python tools/gpu_torch_baseline.py --profile $O/r02a_torch_eager_kernels.md > $O/r02a_gpu_torch_baseline.json 2> $O/r02a_gpu_torch_baseline.err; echo "torch baseline rc=$?"
timeout 300 python tools/profile_elementwise.py > $O/r02a_elementwise.json 2> $O/r02a_elementwise.err; echo "elementwise rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"q_sample|loss_kernel|sampler_step" -s 3 -c 3 -f -o $O/r02a_prof_elementwise python tools/profile_elementwise.py --reps 1 > $O/r02a_ncu_elementwise.log 2>&1; echo "ncu elementwise rc=$?"
for P in tf32 bf16; do
  CMD="python tools/profile_step.py --batch 512 --warm 1 --no-sample --precision $P"
  $CMD > $O/r02a_plain_$P.log 2>&1 || { echo "plain $P failed"; tail -3 $O/r02a_plain_$P.log; continue; }
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:gn_bwd -s 28 -c 4 -f -o $O/r02a_prof_gnbwd_$P $CMD > $O/r02a_ncu_gnbwd_$P.log 2>&1; echo "ncu gn_bwd $P rc=$?"
done
ls -la $O | grep r02a
