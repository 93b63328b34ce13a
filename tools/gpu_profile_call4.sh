#!/bin/bash
# One gpurun call (round 1, fourth profiling pass, after the two-phase tiling / multi-row strip boxes / two-tap weight
# slots / one-instruction dY tile): GPU tests, the default bench, then for each precision a plain run followed by the
# ncu launch list of the same command, then `--set full` captures of the strip and wgrad kernels.
# (B200_PROFILING.md: every ncu run follows a plain run of the same command that exited 0.)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
for P in tf32 bf16; do
  CMD="python tools/profile_step.py --batch 512 --warm 1 --no-sample --precision $P"
  $CMD > gpurun_out/plain_$P.log 2>&1 || { echo "plain $P failed"; tail -3 gpurun_out/plain_$P.log; exit 1; }
  ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches_e_$P.csv $CMD > gpurun_out/ncu_list_e_$P.log 2>&1
done
for P in tf32 bf16; do
  CMD="python tools/profile_step.py --batch 512 --warm 1 --no-sample --precision $P"
  ncu --set full --clock-control none --import-source on -k regex:conv_strip -s 1 -c 3 -f -o gpurun_out/prof_e_strip_$P $CMD > gpurun_out/ncu_e_strip_$P.log 2>&1
  ncu --set full --clock-control none --import-source on -k regex:conv_wgrad -s 32 -c 2 -f -o gpurun_out/prof_e_wgrad_$P $CMD > gpurun_out/ncu_e_wgrad_$P.log 2>&1
done
echo "profile rc=$?"
