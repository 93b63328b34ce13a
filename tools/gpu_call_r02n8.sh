#!/bin/bash
# Round 2: the default bench at N = 8 (torchrun, one rank per GPU), as the driver's scaling pass runs it.
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 20 --warmup 5 > $O/r02_bench_n8.json 2> $O/r02_bench_n8.err; echo "bench n8 rc=$?"; tail -3 $O/r02_bench_n8.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02_bench_n8.json'))
print('ms/step', d['ms_per_step'], 'min', d.get('ms_per_step_min_over_ranks'), 'img/s', d['value'], 'e2e', d['e2e']['value'], 'allreduce_ms', d.get('allreduce_ms'), 'samp', (d.get('sampling') or {}).get('value'), 'bf16', d['bf16']['train_img_s'], d['bf16']['sampling_img_s'], d['clocks'])
PY
