#!/bin/bash
# temporary A/B of the current experiment toggles (1 GPU, same box)
OUT=gpurun_out/exp; mkdir -p $OUT
ms() { grep -o '"ms_per_step": [0-9.]*' | head -1 | cut -d' ' -f2; }
B="python bench.py --steps 300 --warmup 20 --no-e2e --no-tf32 --no-strong --no-loader"
timeout 600 python -m pytest tests/test_gpu_fused_paths.py tests/test_gpu_model.py tests/test_gpu_tf32.py -q -x -m gpu 2>&1 | tail -5
for cfg in "" "B200DP_EXP_WGRAD_DEEP=1" "" "B200DP_EXP_WGRAD_DEEP=1"; do
  v=$(env $cfg timeout 200 $B 2>$OUT/err.log | ms)
  echo "  [${cfg:-default}] ms/step=$v" | tee -a $OUT/ab2.log
done
B200DP_EXP_WGRAD_DEEP=1 timeout 300 python -m pytest tests/test_gpu_kernels.py -q -x -m gpu -k wgrad 2>&1 | tail -2
timeout 100 python scripts/linear_ab.py | tee $OUT/linear_ab2.txt
