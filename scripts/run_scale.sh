#!/bin/bash
# Full measurement set on N GPUs of one box: multi-GPU tests, all-reduce sweep, both bench arms.
#   bash scripts/run_scale.sh N [tests]
N=${1:-2}
OUT=gpurun_out/scale_$N
mkdir -p $OUT
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611"
if [ "$2" = "tests" ]; then
  timeout 900 python -m pytest tests/test_multigpu_native.py -m gpu -q --timeout 600 2>&1 | tail -5 > $OUT/tests.log
  cat $OUT/tests.log
fi
timeout 400 $RUN scripts/allreduce_sweep.py 2>&1 | grep -E "^\{|^#" > $OUT/allreduce_sweep.jsonl
tail -2 $OUT/allreduce_sweep.jsonl
b() { name=$1; shift; timeout 400 $RUN bench.py --gpus $N --steps 100 --warmup 5 "$@" 2>&1 | grep "^{" > $OUT/$name.json; python - <<PY
import json
try:
    r = json.load(open("$OUT/$name.json"))
    print("$name", "ms/step", round(r.get("ms_per_step", 0), 4), "img/s", round(r.get("value", 0)), "e2e", round(r.get("e2e", {}).get("value", 0)), "exposed", r.get("exposed_comm_ms_per_step"))
except Exception as e:
    print("$name FAILED", e)
PY
}
b ref_part3 --impl reference
b native_part3_nvls --nvls on
b native_part3_p2p --nvls off
b native_part2b_nvls --mode part2b --nvls on --no-e2e
b native_part2b_p2p --mode part2b --nvls off --no-e2e
b native_part2a --mode part2a --no-e2e
b native_part2a_extra --mode part2a_extra --no-e2e
b native_part3_strong --scaling strong --no-e2e
b native_part2b_strong --scaling strong --mode part2b --no-e2e
b ref_part3_strong --impl reference --scaling strong --no-e2e
b ref_part2b --impl reference --mode part2b --no-e2e
b ref_part2a --impl reference --mode part2a --no-e2e
