#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $O/r02x_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/r02x_pytest.log
timeout 600 python bench.py --batch 64 --steps 50 --warmup 10 --sample-n 64 --sample-steps 200 --gpu-baseline 0 --small-n 0 --cpu-steps 1 > $O/r02x_bench_b64.json 2> $O/r02x_bench_b64.err; echo "bench b64 rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r02x_bench_b64.json").read().strip().splitlines()[-1])
print("b64 step", d["ms_per_step"], d["value"], "e2e", d["e2e"]["value"], "bf16", d["bf16"]["ms_per_step"], "sampling", d["sampling"]["value"], "guided", d["e2e"]["sampling"]["img_s"])
PY
