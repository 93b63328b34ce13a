#!/bin/bash
# Round 2, two-GPU call: data-parallel parity (overlapped two-bucket allreduce == global-batch step), the bench at N=2
# with the collective between graphs (default) and captured inside the step graph (B200_NCCL_GRAPH=1).
mkdir -p gpurun_out
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29533 tools/dp_check.py > $O/r02f_dp_check.json 2> $O/r02f_dp_check.err; echo "dp_check rc=$?"; cat $O/r02f_dp_check.json; tail -3 $O/r02f_dp_check.err
B200_NCCL_GRAPH=1 timeout 600 $TR --master-port 29534 tools/dp_check.py > $O/r02f_dp_check_graph.json 2> $O/r02f_dp_check_graph.err; echo "dp_check(graph) rc=$?"; cat $O/r02f_dp_check_graph.json; tail -3 $O/r02f_dp_check_graph.err
timeout 900 $TR --master-port 29535 bench.py --gpus 2 --steps 20 --warmup 5 > $O/r02f_bench_n2.json 2> $O/r02f_bench_n2.err; echo "bench n2 rc=$?"
B200_NCCL_GRAPH=1 timeout 900 $TR --master-port 29536 bench.py --gpus 2 --steps 20 --warmup 5 --bf16-leg 0 --sample-n 0 > $O/r02f_bench_n2_ncclgraph.json 2> $O/r02f_bench_n2_ncclgraph.err; echo "bench n2 (nccl in graph) rc=$?"; tail -3 $O/r02f_bench_n2_ncclgraph.err
python - <<'PY'
import json
for f in ('r02f_bench_n2.json','r02f_bench_n2_ncclgraph.json'):
    try:
        d=json.load(open('gpurun_out/'+f))
        print(f, 'ms/step', d['ms_per_step'], 'min', d.get('ms_per_step_min_over_ranks'), 'img/s', d['value'], 'e2e', d['e2e']['value'], 'allreduce_ms', d.get('allreduce_ms'), 'samp', (d.get('sampling') or {}).get('value'))
    except Exception as e:
        print(f, 'unreadable', e)
PY
