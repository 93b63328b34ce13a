#!/bin/bash
# One gpurun call (round 1, third profiling pass, after the strip-epilogue / wgrad two-tap / linear-tile changes):
# for each precision a plain run, then the ncu launch list of the same command; then `--set full` captures of the
# strip kernel (TF32: the bench's roofline kernel; bf16: conv2 with residual + bf16 copy) and of the wgrad kernel.
# (B200_PROFILING.md: every ncu run follows a plain run of the same command that exited 0.)
mkdir -p gpurun_out
for P in tf32 bf16; do
  CMD="python tools/profile_step.py --batch 512 --warm 1 --no-sample --precision $P"
  $CMD > gpurun_out/plain_$P.log 2>&1 || { echo "plain $P failed"; tail -3 gpurun_out/plain_$P.log; exit 1; }
  ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches_c_$P.csv $CMD > gpurun_out/ncu_list_c_$P.log 2>&1
done
for P in tf32 bf16; do
  CMD="python tools/profile_step.py --batch 512 --warm 1 --no-sample --precision $P"
  ncu --set full --clock-control none --import-source on -k regex:conv_strip -s 1 -c 3 -f -o gpurun_out/prof_c_strip_$P $CMD > gpurun_out/ncu_c_strip_$P.log 2>&1
  ncu --set full --clock-control none --import-source on -k regex:conv_wgrad -s 32 -c 2 -f -o gpurun_out/prof_c_wgrad_$P $CMD > gpurun_out/ncu_c_wgrad_$P.log 2>&1
done
echo "profile rc=$?"
