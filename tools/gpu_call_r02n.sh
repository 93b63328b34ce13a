#!/bin/bash
# Programmatic dependent launch (B200_PDL=1) against plain launches: GPU suite under PDL, then the train step at batch
# 64 and 512, sample(16, y) / evaluate() at T = 1000 and the n = 512 sampler.
mkdir -p gpurun_out
O=gpurun_out
B200_PDL=1 timeout 900 python -m pytest tests -m gpu -x -q > $O/r02n_pytest_pdl.log 2>&1; echo "pytest(PDL) rc=$?"; tail -3 $O/r02n_pytest_pdl.log
for P in 0 1; do
  B200_PDL=$P timeout 600 python bench.py --batch 64 --steps 50 --warmup 10 --sample-n 64 --sample-steps 200 --gpu-baseline 0 --small-n 1 --cpu-steps 1 > $O/r02n_bench_b64_pdl$P.json 2> $O/r02n_bench_b64_pdl$P.err; echo "b64 pdl=$P rc=$?"
  B200_PDL=$P timeout 600 python bench.py --steps 20 --warmup 5 --sample-steps 200 --gpu-baseline 0 --small-n 0 --cpu-steps 1 > $O/r02n_bench_b512_pdl$P.json 2> $O/r02n_bench_b512_pdl$P.err; echo "b512 pdl=$P rc=$?"
done
python - <<'PY'
import json
for b in (64, 512):
    for p in (0, 1):
        try:
            d = json.loads(open(f"gpurun_out/r02n_bench_b{b}_pdl{p}.json").read().strip().splitlines()[-1])
        except Exception as e:
            print(b, p, "unreadable", e); continue
        sm = d.get("sampling_small_n") or {}
        print(f"B={b} PDL={p}: step {d['ms_per_step']:.3f} ms, e2e {d['e2e']['ms_per_step']:.3f}, bf16 {d['bf16']['ms_per_step']:.3f}, "
              f"sampling {d['sampling']['value']:.1f} img/s, guided {d['e2e']['sampling']['img_s']:.1f}, "
              f"sample16 {(sm.get('sample16_guided') or {}).get('ms_per_step')}, eval25 {(sm.get('evaluate25') or {}).get('ms_per_step')}, clocks {d['clocks']['sm_mhz']}")
PY
