#!/bin/bash
# GroupNorm with the specialised phase-2 loops / running pointers: tests, per-kernel times, the default bench
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $O/r02q_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/r02q_pytest.log
timeout 300 python tools/time_gn.py > $O/r02q_time_gn.txt 2>&1
timeout 300 python tools/time_gn.py --bf16 1 >> $O/r02q_time_gn.txt 2>&1
cat $O/r02q_time_gn.txt
timeout 900 python bench.py --steps 20 --warmup 5 --gpu-baseline 0 --cpu-steps 1 > $O/r02q_bench_n1.json 2> $O/r02q_bench_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r02q_bench_n1.json").read().strip().splitlines()[-1])
print("step", d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], "bf16", d["bf16"]["ms_per_step"], "sampling", d["sampling"]["value"],
      "bf16 sampling", d["bf16"]["sampling_img_s"], "clocks", d["clocks"])
for k in ("b200_gn_bwd", "b200_gn_fwd"):
    print(k, d["breakdown"][k], d["bf16"]["breakdown"][k])
PY
