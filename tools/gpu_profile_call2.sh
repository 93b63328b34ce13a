#!/bin/bash
# One gpurun call (round 1, second profiling pass): for each precision a plain run, the ncu launch list of the same
# command, then `--set full` captures of the bf16 strip / wgrad / GroupNorm kernels at the 28x28 level.
# (B200_PROFILING.md: every ncu run follows a plain run of the same command that exited 0.)
mkdir -p gpurun_out
for P in tf32 bf16; do
  CMD="python tools/profile_step.py --batch 512 --warm 1 --no-sample --precision $P"
  $CMD > gpurun_out/plain_$P.log 2>&1 || { echo "plain $P failed"; tail -3 gpurun_out/plain_$P.log; exit 1; }
  ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches_$P.csv $CMD > gpurun_out/ncu_list_$P.log 2>&1
done
CMD="python tools/profile_step.py --batch 512 --warm 1 --no-sample --precision bf16"
ncu --set full --clock-control none --import-source on -k regex:conv_strip -s 1 -c 3 -f -o gpurun_out/prof_strip_bf16 $CMD > gpurun_out/ncu_strip_bf16.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:conv_wgrad -s 32 -c 3 -f -o gpurun_out/prof_wgrad_bf16 $CMD > gpurun_out/ncu_wgrad_bf16.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gn_ -s 50 -c 4 -f -o gpurun_out/prof_gn_bf16 $CMD > gpurun_out/ncu_gn_bf16.log 2>&1
echo "profile rc=$?"
