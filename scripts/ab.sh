#!/bin/bash
# A/B a compile-time knob on the GPU box: scripts/ab.sh "-DEVAL_MINB=4" ["-D..."] ; prints ms/step + breakdown per variant
for extra in "$@"; do
  BSNLS_NVCC_EXTRA="$extra" python beam_slam_b200/build.py --force > /dev/null 2>&1 || { echo "build failed: $extra"; continue; }
  echo "== $extra"
  python bench.py --steps 20 --warmup 5 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],4), {k:round(v,4) for k,v in d['kernel_ms'].items()})"
done
