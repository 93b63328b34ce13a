#!/bin/bash
# staged (bulk-copy) strip epilogue: parity, per-layer timing with and without, short bench
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_convs.py tests/test_gpu_bf16.py -q -x > $O/r02j_pytest_convs.log 2>&1; echo "conv tests rc=$?"; tail -6 $O/r02j_pytest_convs.log
for S in 1 0; do
  echo "== B200_STRIP_STAGE=$S"
  B200_STRIP_STAGE=$S timeout 300 python tools/time_conv.py 2>&1 | grep -E "mode=[01] (res=[01]|bf16)"
done > $O/r02j_stage.txt 2>&1
cat $O/r02j_stage.txt
timeout 600 python bench.py --steps 20 --warmup 5 --gpu-baseline 0 --small-n 0 --sample-api 0 --cpu-steps 1 > $O/r02j_bench_n1.json 2> $O/r02j_bench_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02j_bench_n1.json'))
print('ms/step', d['ms_per_step'], 'img/s', d['value'], 'samp', d['sampling']['value'], 'bf16', d['bf16']['ms_per_step'], d['bf16']['sampling_img_s'], d['clocks'])
for k,v in list(d['breakdown'].items())[:5]: print('  ',k,v)
for k,v in list(d['bf16']['breakdown'].items())[:5]: print(' bf16 ',k,v)
PY
