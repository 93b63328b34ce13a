#!/usr/bin/env bash
# Install the UNMODIFIED reference (matwilso/generative_models, python package `gms`) into the git-ignored
# baseline/_ref/ so that it travels to the GPU box with the gpurun snapshot (it is git-ignored, not gpurun-ignored).
# Used only as the comparison arm: bench.py --impl reference / cpu_baseline (oracle/cpu_reference_bench.py),
# tools/gpu_torch_baseline.py, tools/run_gms_main.py --reference-root baseline/_ref, and the -m gpu drop-in test.
# /root/reference is read-only, so the install runs from a copy under /tmp; --no-deps: its requirements.txt pins
# nothing that is missing here except ignite / matplotlib / moviepy, which are stubbed or optional.
set -euo pipefail
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
SRC="${GMS_REFERENCE:-/root/reference}"
DST="$ROOT/baseline/_ref"
if [ ! -f "$SRC/setup.py" ]; then
  echo "vendor_reference: $SRC not present (GPU box?) - keeping $DST as is" >&2
  exit 0
fi
TMP="$(mktemp -d)"
cp -r "$SRC" "$TMP/ref"
rm -rf "$DST"
python -m pip install -q --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse --target "$DST" "$TMP/ref"
rm -rf "$TMP"
test -f "$DST/gms/diffusion/diffusion_model.py"
echo "vendor_reference: installed $(cd "$DST" && find gms -name '*.py' | wc -l) python files into $DST"
