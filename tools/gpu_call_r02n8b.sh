#!/bin/bash
# Round 2, eight GPUs: BASELINE.json configs[2] (bf16, global batch 4096 = 512/GPU) and configs[3] (65 536 images x 1000
# steps, 8 192 per GPU, no communication) as stated.
mkdir -p gpurun_out
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29542 bench.py --gpus 8 --config 2 --steps 20 --warmup 5 > $O/r02_bench_config2_n8.json 2> $O/r02_bench_config2_n8.err; echo "config2 n8 rc=$?"; tail -2 $O/r02_bench_config2_n8.err
timeout 500 $TR --master-port 29543 bench.py --gpus 8 --config 3 --precision tf32 > $O/r02_bench_config3_65536_n8_tf32.json 2> $O/r02_bench_config3_65536_n8_tf32.err; echo "config3 n8 rc=$?"; tail -2 $O/r02_bench_config3_65536_n8_tf32.err
python - <<'PY'
import json
for f in ('r02_bench_config2_n8.json','r02_bench_config3_65536_n8_tf32.json'):
    try:
        d=json.load(open('gpurun_out/'+f)); print(f, d['value'], d['ms_per_step'], d.get('seconds'), d['config']['workload'])
    except Exception as e: print(f, 'unreadable', e)
PY
