#!/bin/bash
# One gpurun call that validates and A/Bs everything in DESIGN.md section 8 on ONE GPU:
#   /usr/local/graft/bin/gpurun --timeout 900 -- 'bash scripts/validate_experimental.sh'
# then, with 2 GPUs, the multi-rank items:
#   /usr/local/graft/bin/gpurun --gpus 2 --timeout 900 -- 'bash scripts/validate_experimental.sh multi'
# Results land in gpurun_out/experimental/.  Same box, same clocks for every A/B pair.
OUT=gpurun_out/experimental
mkdir -p $OUT
ms() { grep -o '"ms_per_step": [0-9.]*' | head -1 | cut -d' ' -f2; }

if [ "$1" != "multi" ]; then
  echo "== acceptance tests (single GPU)"
  B200DP_EXPERIMENTAL=1 timeout 600 python -m pytest tests/test_gpu_experimental.py -m "gpu and not multigpu" -q \
      --timeout 300 2>&1 | tail -15 | tee $OUT/tests_1gpu.log
  echo "== the regular single-GPU suite with the bulk-copy BatchNorm kernels behind the plain entry points"
  B200DP_BN_TMA=1 timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_model.py -m gpu -q \
      --timeout 300 2>&1 | tail -3 | tee $OUT/tests_bn_tma.log
  echo "== A/B, 1 GPU, 100 steps each"
  for cfg in "" "B200DP_DGRAD_BN_REDUCE=1" "B200DP_HEAD_FUSED=1" "B200DP_PERSIST=1" "B200DP_ZERO_PER_LAYER=1" "B200DP_BN_TMA=1" \
             "B200DP_DGRAD_BN_REDUCE=1 B200DP_HEAD_FUSED=1 B200DP_PERSIST=1 B200DP_ZERO_PER_LAYER=1 B200DP_BN_TMA=1" ""; do
    v=$(env $cfg timeout 200 python bench.py --steps 100 --no-e2e 2>&1 | ms)
    echo "  [${cfg:-baseline}] ms/step=$v" | tee -a $OUT/ab_1gpu.log
  done
else
  N=$(nvidia-smi -L | wc -l)
  RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29631"
  echo "== acceptance test ($N GPUs)"
  B200DP_EXPERIMENTAL=1 timeout 600 python -m pytest tests/test_gpu_experimental.py -m "gpu and multigpu" -q \
      --timeout 400 2>&1 | tail -15 | tee $OUT/tests_${N}gpu.log
  echo "== A/B, $N GPUs part3, 100 steps each"
  for cfg in "" "B200DP_LAZY_EXIT_BARRIER=1" "B200DP_PART3_PREP_OVERLAP=1" "B200DP_WGRAD_PUSH=1" \
             "B200DP_LAZY_EXIT_BARRIER=1 B200DP_WGRAD_PUSH=1" ""; do
    v=$(env $cfg timeout 200 $RUN bench.py --gpus $N --steps 100 --no-e2e 2>&1 | ms)
    echo "  [${cfg:-baseline}] ms/step=$v" | tee -a $OUT/ab_${N}gpu.log
  done
  echo "== bucket plans"
  for bp in layer cap:9 cap:12 ddp single; do
    v=$(timeout 200 $RUN bench.py --gpus $N --steps 100 --no-e2e --bucket-plan $bp 2>&1 | ms)
    echo "  [bucket-plan $bp] ms/step=$v" | tee -a $OUT/ab_${N}gpu.log
  done
  echo "== reducer CTAs"
  for cb in 16 24 32 48 64; do
    v=$(timeout 200 $RUN bench.py --gpus $N --steps 100 --no-e2e --comm-blocks $cb 2>&1 | ms)
    echo "  [comm-blocks $cb] ms/step=$v" | tee -a $OUT/ab_${N}gpu.log
  done
fi
