#!/bin/bash
# One gpurun call: plain run, ncu launch list of the same command, then `--set full` captures of the top kernels.
# (B200_PROFILING.md: every ncu run follows a plain run of the same command that exited 0.)
mkdir -p gpurun_out
CMD="python tools/profile_step.py --batch 512 --warm 1 --no-sample"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:conv_strip -s 30 -c 3 -f -o gpurun_out/prof_strip $CMD > gpurun_out/ncu_strip.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:conv_wgrad -s 36 -c 3 -f -o gpurun_out/prof_wgrad $CMD > gpurun_out/ncu_wgrad.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gn_ -s 50 -c 4 -f -o gpurun_out/prof_gn $CMD > gpurun_out/ncu_gn.log 2>&1
echo "profile rc=$?"
tail -2 gpurun_out/plain.log
