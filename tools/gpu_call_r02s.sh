#!/bin/bash
# Strip epilogue with one-instruction row addresses: conv / bf16 / model tests, per-layer times, default bench
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $O/r02s_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/r02s_pytest.log
timeout 300 python tools/time_conv.py > $O/r02s_time_conv.txt 2>&1; head -32 $O/r02s_time_conv.txt
timeout 900 python bench.py --steps 20 --warmup 5 --gpu-baseline 0 --cpu-steps 1 > $O/r02s_bench_n1.json 2> $O/r02s_bench_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r02s_bench_n1.json").read().strip().splitlines()[-1])
print("step", d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], "bf16", d["bf16"]["ms_per_step"], "sampling", d["sampling"]["value"],
      "guided", d["e2e"]["sampling"]["img_s"], "bf16 sampling", d["bf16"]["sampling_img_s"], "clocks", d["clocks"], "roofline", d["roofline"]["frac"], d["bf16"]["roofline"]["frac"])
for k, v in list(d["breakdown"].items())[:5]: print(k, v)
for k, v in list(d["bf16"]["breakdown"].items())[:5]: print("bf16", k, v)
PY
