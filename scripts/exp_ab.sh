#!/bin/bash
# temporary: validation + A/B of the weight re-layout behind the buckets (2 GPUs)
OUT=gpurun_out/exp; mkdir -p $OUT
N=2
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611"
timeout 900 python -m pytest tests/test_multigpu_native.py tests/test_gpu_fused_paths.py tests/test_gpu_model.py -q -x -m "gpu" --timeout 600 2>&1 | tail -4
ms() { python -c "import json,sys; r=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); print(round(r['ms_per_step'],4), round(r.get('exposed_comm_fraction',0),4))"; }
v=$(CUDA_VISIBLE_DEVICES=0 timeout 200 python bench.py --steps 300 --warmup 20 --no-e2e --no-tf32 --no-strong --no-loader 2>$OUT/err.log | ms); echo "  [1 GPU] $v" | tee -a $OUT/ab5.log
ab() { v=$(timeout 300 $RUN bench.py --gpus $N --steps 200 --warmup 20 --no-e2e --no-strong --no-tf32 --no-selfcheck --no-loader "$@" 2>&1 | ms); echo "  [$LABEL$*] $v" | tee -a $OUT/ab5.log; }
ab
LABEL="prep-at-tail " B200DP_EXP_PREP_TAIL=1 ab
ab
LABEL="prep-at-tail " B200DP_EXP_PREP_TAIL=1 ab
ab --bucket-plan cap:12+last
ab --bucket-plan cap:16
LABEL="prep-at-tail " B200DP_EXP_PREP_TAIL=1 ab --bucket-plan cap:12+last
timeout 300 $RUN scripts/timeline_step.py --mode part3 --out $OUT/timeline_part3_2gpu > $OUT/timeline_part3_2gpu.log 2>&1; rm -f $OUT/timeline_part3_2gpu.json
