#!/bin/bash
# batch 64 / 128 / 256: strip-vs-im2col crossover (B200_STRIP_MIN_TILES, default 296 = 2 tiles per SM)
mkdir -p gpurun_out
O=gpurun_out
rm -f $O/r02w.txt
for B in 64 128 256; do
  for MT in 296 148 74 0; do
    B200_STRIP_MIN_TILES=$MT timeout 300 python bench.py --batch $B --steps 40 --warmup 10 --sample-n $B --sample-steps 100 --gpu-baseline 0 --small-n 0 --cpu-steps 1 --bf16-leg 0 --sample-api 0 > $O/r02w_tmp.json 2> $O/r02w_tmp.err
    python - "$B" "$MT" >> $O/r02w.txt <<'PY'
import json, sys
d = json.loads(open("gpurun_out/r02w_tmp.json").read().strip().splitlines()[-1])
print(f"B={sys.argv[1]} MIN_TILES={sys.argv[2]}: step {d['ms_per_step']:.3f} ms  {d['value']:.0f} img/s  sampling {d['sampling']['value']:.1f} img/s (T={d['sampling']['timesteps']})")
PY
  done
done
cat $O/r02w.txt
