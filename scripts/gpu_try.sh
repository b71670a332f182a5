#!/usr/bin/env bash
# experiment round: parity tests, bench + launch list, sort variants, unique tile sizes
set -u
mkdir -p gpurun_out
bash scripts/gpu_quick.sh
echo "--- sort variants"
VARIANTS=${VARIANTS:-20,30,31,32,33,34} timeout 600 python scripts/sort_bench.py 2>&1 | tee gpurun_out/sort_variants.txt
for kb in 16 64; do
  echo "--- MTG_UB_TILE_KB=$kb"
  MTG_UB_TILE_KB=$kb timeout 300 python bench.py --steps 3 --warmup 2 --no-e2e --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['phases_ms'])"
done
