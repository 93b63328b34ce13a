#!/bin/bash
# Round 2, fifth call: GroupNorm backward with running pointers — GPU suite, stand-alone timings, per-op list, short bench.
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > $O/r02e_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 $O/r02e_pytest_gpu.log
timeout 300 python tools/time_gn.py > $O/r02e_time_gn.txt 2>&1; echo "time_gn rc=$?"; cat $O/r02e_time_gn.txt
timeout 300 python tools/per_op.py --csv $O/r02e_per_op_tf32.csv > /dev/null 2>&1; echo "per_op rc=$?"
timeout 600 python bench.py --steps 20 --warmup 5 --gpu-baseline 0 --small-n 0 --sample-api 0 --cpu-steps 2 > $O/r02e_bench_n1.json 2> $O/r02e_bench_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02e_bench_n1.json'))
print('ms/step', d['ms_per_step'], 'img/s', d['value'], 'e2e', d['e2e']['value'], 'samp', d['sampling']['value'], 'bf16', d['bf16']['ms_per_step'], d['bf16']['sampling_img_s'])
for k,v in list(d['breakdown'].items())[:6]: print('  ',k,v)
for k,v in list(d['bf16']['breakdown'].items())[:6]: print(' bf16 ',k,v)
PY
