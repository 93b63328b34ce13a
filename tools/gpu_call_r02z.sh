#!/bin/bash
# Round 2, final measurement call (one B200): GPU suite, the default bench (as the driver runs it) and the reference arm,
# the unmodified gms.main at full size, the other BASELINE.json configs, then the ncu evidence — every ncu run follows a
# plain run of the same command that exited 0 (B200_PROFILING.md).
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > $O/r02z_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $O/r02z_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/r02z_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/r02z_smoke.log
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/r02z_bench_n1.json 2> $O/r02z_bench_n1.err; echo "bench rc=$?"; tail -2 $O/r02z_bench_n1.err
timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > $O/r02z_bench_reference.json 2> $O/r02z_bench_reference.err; echo "reference arm rc=$?"
timeout 900 python tools/run_gms_main.py --synthetic-mnist --git-cwd -- --model diffusion_model --bs 512 --epochs 1 --mean_type eps --sampler noisy --timesteps 1000 --eval_heavy 0 --logdir /tmp/gms_logs > $O/r02z_gms_main_b512_T1000.log 2>&1; echo "gms.main rc=$?"; tail -4 $O/r02z_gms_main_b512_T1000.log
timeout 900 python bench.py --config 2 --steps 10 --warmup 3 > $O/r02z_bench_config2_n1.json 2> $O/r02z_bench_config2_n1.err; echo "config2 rc=$?"; tail -2 $O/r02z_bench_config2_n1.err
timeout 900 python bench.py --config 3 --config3-images 8192 --precision tf32 > $O/r02z_bench_config3_8192_tf32.json 2> $O/r02z_bench_config3_8192_tf32.err; echo "config3 tf32 rc=$?"
timeout 900 python bench.py --config 3 --config3-images 8192 --precision bf16 > $O/r02z_bench_config3_8192_bf16.json 2> $O/r02z_bench_config3_8192_bf16.err; echo "config3 bf16 rc=$?"
timeout 900 python bench.py --config 4 --steps 10 --warmup 3 --sample-steps 100 > $O/r02z_bench_config4_n1.json 2> $O/r02z_bench_config4_n1.err; echo "config4 rc=$?"; tail -2 $O/r02z_bench_config4_n1.err
timeout 300 python tools/profile_elementwise.py > $O/r02z_elementwise.json 2> $O/r02z_elementwise.err; echo "elementwise rc=$?"
timeout 300 python tools/time_gn.py > $O/r02z_time_gn.txt 2>&1
timeout 300 python tools/time_conv.py > $O/r02z_time_conv.txt 2>&1
# ---- ncu: launch list of the bench command itself (short form), then --set full captures
BCMD="python bench.py --steps 2 --warmup 3 --sample-steps 3 --bf16-leg 0 --gpu-baseline 0 --small-n 0 --sample-api 0 --cpu-steps 1"
$BCMD > $O/r02z_bench_short.json 2> $O/r02z_bench_short.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/r02z_launches_bench_tf32.csv $BCMD > $O/r02z_ncu_list_bench.log 2>&1; echo "ncu list rc=$?"
for P in tf32 bf16; do
  CMD="python tools/profile_step.py --batch 512 --warm 1 --no-sample --precision $P"
  $CMD > $O/r02z_plain_$P.log 2>&1 || { echo "plain $P failed"; continue; }
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_strip -s 1 -c 3 -f -o $O/r02z_prof_strip_$P $CMD > $O/r02z_ncu_strip_$P.log 2>&1
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_wgrad -s 32 -c 2 -f -o $O/r02z_prof_wgrad_$P $CMD > $O/r02z_ncu_wgrad_$P.log 2>&1
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:gn_bwd -s 28 -c 4 -f -o $O/r02z_prof_gnbwd_$P $CMD > $O/r02z_ncu_gnbwd_$P.log 2>&1
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:gn_fwd -s 26 -c 2 -f -o $O/r02z_prof_gnfwd_$P $CMD > $O/r02z_ncu_gnfwd_$P.log 2>&1
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"q_sample|loss_|sampler_step" -s 3 -c 3 -f -o $O/r02z_prof_elementwise python tools/profile_elementwise.py --reps 1 > $O/r02z_ncu_elementwise.log 2>&1; echo "ncu elementwise rc=$?"
ls -la $O | grep r02z | wc -l
