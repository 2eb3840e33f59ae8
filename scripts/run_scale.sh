#!/bin/bash
# Measurement set on N GPUs of one box: (optionally) the multi-GPU tests, both bench arms with every extra key,
# then short A/B runs of the part3 knobs.   bash scripts/run_scale.sh N [tests]
N=${1:-2}
OUT=gpurun_out/scale_$N
mkdir -p $OUT
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611"
if [ "$2" = "tests" ]; then
  timeout 1500 python -m pytest tests/test_multigpu_native.py -m gpu -q --timeout 900 2>&1 | tail -15 > $OUT/tests.log
  cat $OUT/tests.log
elif [ "$2" = "sometests" ]; then      # the expensive world sizes: what the last kernel changes can reach
  timeout 900 python -m pytest tests/test_multigpu_native.py tests/test_gpu_fused_paths.py -m "gpu and multigpu" -q --timeout 600 \
      -k "push or resume or selfcheck or synced_gradient" 2>&1 | tail -15 > $OUT/tests.log
  cat $OUT/tests.log
fi
timeout 600 $RUN bench.py --gpus $N --steps 100 --warmup 5 2>&1 | grep "^{" > $OUT/native.json
timeout 600 $RUN bench.py --gpus $N --steps 100 --warmup 5 --impl reference 2>&1 | grep "^{" > $OUT/reference.json
python - <<PY
import json
for name in ("native", "reference"):
    try:
        r = json.load(open("$OUT/%s.json" % name))
        print(name, "ms/step", round(r["ms_per_step"], 4), "img/s", round(r["value"]), "e2e", round(r.get("e2e", {}).get("value", 0)),
              "exposed", r.get("exposed_comm_fraction"), "tf32", (r.get("same_precision") or {}).get("ms_per_step"),
              "strong", {k: round(v["ms_per_step"], 4) for k, v in r.get("strong", {}).items() if isinstance(v, dict) and "ms_per_step" in v},
              "selfcheck", r.get("selfcheck"))
    except Exception as e:
        print(name, "FAILED", e)
PY
timeout 300 $RUN scripts/timeline_step.py --mode part3 --out $OUT/timeline_part3 > $OUT/timeline_part3.log 2>&1
rm -f $OUT/timeline_part3.json
ab() { v=$(timeout 300 $RUN bench.py --gpus $N --steps 100 --no-e2e --no-strong --no-tf32 --no-selfcheck --no-loader "$@" 2>&1 | grep "^{" | python -c "import json,sys; r=json.loads(sys.stdin.read()); print(round(r['ms_per_step'],4), round(r.get('exposed_comm_fraction',0),4))"); echo "  [${LABEL}$*] $v" | tee -a $OUT/ab.log; }
ab
LABEL="re-layout at the tail, lazy exit barriers " B200DP_EXP_PREP_TAIL=1 ab
ab --bucket-plan cap:16+last
ab --bucket-plan cap:12
ab
LABEL="re-layout at the tail, lazy exit barriers " B200DP_EXP_PREP_TAIL=1 ab
timeout 300 $RUN bench.py --gpus $N --steps 100 --no-strong --no-tf32 --no-selfcheck --loader-threads 8 2>&1 | grep "^{" > $OUT/native_loader8.json
python -c "import json; r=json.load(open('$OUT/native_loader8.json')); print('loader threads 8:', r.get('e2e_with_loader'))"
if [ "$3" = "modes" ]; then
  ab --mode part2b
  ab --mode part2a
  ab --mode part2a_extra
fi
