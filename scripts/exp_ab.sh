#!/bin/bash
# temporary: validation + bench of the current build (1 GPU)
OUT=gpurun_out/exp; mkdir -p $OUT
ms() { grep -o '"ms_per_step": [0-9.]*' | head -1 | cut -d' ' -f2; }
B="python bench.py --steps 300 --warmup 20 --no-e2e --no-tf32 --no-strong --no-loader"
timeout 900 python -m pytest tests -q -x -m "gpu and not multigpu" 2>&1 | tail -4
for cfg in "" ""; do
  v=$(env $cfg timeout 200 $B 2>$OUT/err.log | ms)
  echo "  [${cfg:-default}] ms/step=$v" | tee -a $OUT/ab3.log
done
v=$(timeout 200 $B --dtype tf32 2>$OUT/err.log | ms); echo "  [tf32] ms/step=$v" | tee -a $OUT/ab3.log
timeout 300 python scripts/timeline_step.py --mode part1 --out $OUT/timeline_part1 > $OUT/timeline_part1.log 2>&1; rm -f $OUT/timeline_part1.json
