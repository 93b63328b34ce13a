#!/bin/bash
# Round 2, fourth call: GroupNorm reductions without shared-memory atomics, 5 strip buffers for the 1x1 layers, the
# warp-per-sample loss kernel — GPU suite, stand-alone timings, a short bench.
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > $O/r02d_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -25 $O/r02d_pytest_gpu.log
timeout 300 python tools/time_gn.py > $O/r02d_time_gn.txt 2>&1; echo "time_gn rc=$?"; cat $O/r02d_time_gn.txt
timeout 300 python tools/time_conv.py 2>&1 | grep -E "R=1|wgrad" > $O/r02d_time_conv.txt; cat $O/r02d_time_conv.txt
timeout 300 python tools/profile_elementwise.py > $O/r02d_elementwise.json 2> $O/r02d_elementwise.err; echo "elementwise rc=$?"; cat $O/r02d_elementwise.json
timeout 600 python bench.py --steps 20 --warmup 5 --gpu-baseline 0 --small-n 0 --sample-api 0 --cpu-steps 2 > $O/r02d_bench_n1.json 2> $O/r02d_bench_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02d_bench_n1.json'))
print('ms/step', d['ms_per_step'], 'img/s', d['value'], 'e2e', d['e2e']['value'], 'samp', d['sampling']['value'], 'bf16', d['bf16']['ms_per_step'], d['bf16']['sampling_img_s'])
for k,v in d['breakdown'].items(): print('  ',k,v)
for k,v in d['bf16']['breakdown'].items(): print(' bf16 ',k,v)
PY
