#!/bin/bash
# temporary A/B of the current experiment toggles (1 GPU, same box)
OUT=gpurun_out/exp; mkdir -p $OUT
ms() { grep -o '"ms_per_step": [0-9.]*' | head -1 | cut -d' ' -f2; }
B="python bench.py --steps 300 --warmup 20 --no-e2e --no-tf32 --no-strong --no-loader"
timeout 300 python -m pytest tests/test_gpu_model.py -q -x -m gpu 2>&1 | tail -2
for cfg in "" "B200DP_EXP_FILL=head" "B200DP_EXP_PRIO=1" "B200DP_EXP_SPLIT0=2" "B200DP_EXP_SPLIT0=4" "B200DP_EXP_SPLIT0=2 B200DP_EXP_PRIO=1" ""; do
  v=$(env $cfg timeout 200 $B 2>$OUT/err.log | ms)
  echo "  [${cfg:-default}] ms/step=$v" | tee -a $OUT/ab.log
done
B200DP_EXP_SPLIT0=2 timeout 300 python -m pytest tests/test_gpu_model.py -q -x -m gpu 2>&1 | tail -2
timeout 100 python scripts/linear_ab.py | tee $OUT/linear_ab.txt
timeout 300 python bench.py --steps 100 --no-strong --no-tf32 > $OUT/bench_loader_native.json 2>$OUT/bench_loader_native.err; tail -c 600 $OUT/bench_loader_native.json
timeout 300 python bench.py --impl reference --steps 100 --no-strong > $OUT/bench_loader_reference.json 2>$OUT/bench_loader_reference.err; tail -c 500 $OUT/bench_loader_reference.json; tail -3 $OUT/bench_loader_reference.err
