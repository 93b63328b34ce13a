#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
rm -f $O/r02w2.txt
for MT in 296 0 296 0; do
  B200_STRIP_MIN_TILES=$MT timeout 400 python bench.py --steps 20 --warmup 5 --sample-steps 100 --gpu-baseline 0 --small-n 1 --cpu-steps 1 --bf16-leg 0 --sample-api 0 > $O/r02w_tmp.json 2> $O/r02w_tmp.err
  python - "$MT" >> $O/r02w2.txt <<'PY'
import json, sys
d = json.loads(open("gpurun_out/r02w_tmp.json").read().strip().splitlines()[-1])
sm = d.get("sampling_small_n") or {}
print(f"B=512 MIN_TILES={sys.argv[1]}: step {d['ms_per_step']:.3f} ms  sampling {d['sampling']['value']:.1f} img/s  sample16 {sm['sample16_guided']['ms_per_step']:.4f} ms/step  eval25 {sm['evaluate25']['ms_per_step']:.4f}  clocks {d['clocks']['sm_mhz']}")
PY
done
cat $O/r02w2.txt
