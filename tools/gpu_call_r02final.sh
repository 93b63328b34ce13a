#!/bin/bash
# Round 2, closing measurement call (one B200) with the final code: GPU suite, smoke, the default bench exactly as the
# driver runs it, the reference arm, the ncu launch list of the bench command and --set full captures of the top kernels
# (every ncu run follows a plain run of the same command that exited 0 — B200_PROFILING.md).
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -q > $O/r02f_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 $O/r02f_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/r02f_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/r02f_smoke.log
timeout 900 python bench.py > $O/r02f_bench_n1.json 2> $O/r02f_bench_n1.err; echo "bench rc=$?"; tail -2 $O/r02f_bench_n1.err
timeout 600 python bench.py --impl reference > $O/r02f_bench_reference.json 2> $O/r02f_bench_reference.err; echo "reference arm rc=$?"
timeout 300 python tools/time_conv.py > $O/r02f_time_conv.txt 2>&1
BCMD="python bench.py --steps 2 --warmup 3 --sample-steps 3 --bf16-leg 0 --gpu-baseline 0 --small-n 0 --sample-api 0 --cpu-steps 1"
$BCMD > $O/r02f_bench_short.json 2> $O/r02f_bench_short.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/r02f_launches_bench_tf32.csv $BCMD > $O/r02f_ncu_list_bench.log 2>&1; echo "ncu list rc=$?"
for P in tf32 bf16; do
  CMD="python tools/profile_step.py --batch 512 --warm 1 --no-sample --precision $P"
  $CMD > $O/r02f_plain_$P.log 2>&1 || { echo "plain $P failed"; continue; }
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_strip -s 1 -c 3 -f -o $O/r02f_prof_strip_$P $CMD > $O/r02f_ncu_strip_$P.log 2>&1
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:gn_bwd -s 28 -c 4 -f -o $O/r02f_prof_gnbwd_$P $CMD > $O/r02f_ncu_gnbwd_$P.log 2>&1
done
ls $O | grep r02f | wc -l
