#!/bin/bash
# temporary A/B: BatchNorm pipelines with 3 CTAs per SM (1 GPU)
OUT=gpurun_out/exp; mkdir -p $OUT
ms() { grep -o '"ms_per_step": [0-9.]*' | head -1 | cut -d' ' -f2; }
B="python bench.py --steps 300 --warmup 20 --no-e2e --no-tf32 --no-strong --no-loader"
B200DP_EXP_BN_CTAS=3 B200DP_EXP_BN_POOL3=1 timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_model.py tests/test_gpu_tf32.py -q -x -m gpu 2>&1 | tail -3
timeout 600 python -m pytest tests -q -x -m "gpu and not multigpu" 2>&1 | tail -3
for cfg in "" "B200DP_EXP_BN_CTAS=3" "B200DP_EXP_BN_POOL3=1" "B200DP_EXP_BN_CTAS=3 B200DP_EXP_BN_POOL3=1" "" "B200DP_EXP_BN_CTAS=3 B200DP_EXP_BN_POOL3=1"; do
  v=$(env $cfg timeout 200 $B 2>$OUT/err.log | ms)
  echo "  [${cfg:-default}] ms/step=$v" | tee -a $OUT/ab6.log
done
