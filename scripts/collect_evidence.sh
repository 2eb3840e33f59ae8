#!/bin/bash
# One single-GPU gpurun call that produces the round's one-GPU evidence:
#   /usr/local/graft/bin/gpurun --timeout 2400 -- 'bash scripts/collect_evidence.sh [notests]'
# 1. the GPU test suite (single-GPU tests)      -> gpurun_out/evidence/tests_1gpu.log
# 2. bench.py, both arms (+ the tf32 arm)        -> gpurun_out/evidence/bench_1gpu_{native,reference}.json
# 3. the in-graph timeline of the replayed step  -> gpurun_out/evidence/timeline_part1.txt
# 4. ONE ncu run (--set full, source imported) over a fixed choice of kernels of one eager step, after the same
#    command has exited 0 without ncu             -> gpurun_out/evidence/step.ncu-rep (+ raw page as csv)
OUT=gpurun_out/evidence
mkdir -p $OUT
if [ "$1" != "notests" ]; then
  timeout 1500 python -m pytest tests -m "gpu and not multigpu" -q --timeout 600 > $OUT/tests_1gpu.log 2>&1
  tail -3 $OUT/tests_1gpu.log
fi
timeout 400 python bench.py --impl reference --steps 100 --warmup 10 > $OUT/bench_1gpu_reference.json 2> $OUT/bench_1gpu_reference.err
timeout 600 python bench.py --steps 200 --warmup 20 > $OUT/bench_1gpu_native.json 2> $OUT/bench_1gpu_native.err
python - <<'EOF'
import json
for arm in ("native", "reference"):
    try:
        d = json.loads(open(f"gpurun_out/evidence/bench_1gpu_{arm}.json").read().strip().splitlines()[-1])
        print(arm, "ms/step", d.get("ms_per_step"), "img/s", d.get("value"), "e2e", (d.get("e2e") or {}).get("value"),
              "launches", d.get("gpu_launches"), "tf32", (d.get("same_precision") or {}).get("ms_per_step"),
              "clocks", d.get("clocks"))
    except Exception as e:
        print(arm, "no result:", e)
EOF
timeout 300 python scripts/timeline_step.py --mode part1 --out $OUT/timeline_part1 > $OUT/timeline_part1.log 2>&1
tail -4 $OUT/timeline_part1.log
rm -f $OUT/timeline_part1.json
if timeout 300 python scripts/profile_step.py --select > $OUT/profile_step.log 2>&1; then
  timeout 1500 ncu --set full --clock-control none --import-source on --profile-from-start off -f -o $OUT/step \
      python scripts/profile_step.py --select > $OUT/ncu.log 2>&1
  tail -3 $OUT/ncu.log
  ncu -i $OUT/step.ncu-rep --page raw --csv > $OUT/step_raw.csv 2> /dev/null
  ls -la $OUT
  # what travels back is capped at 64 MiB: the csv is the fallback if the report is too large
  if [ $(stat -c %s $OUT/step.ncu-rep) -gt 57000000 ]; then
    ncu -i $OUT/step.ncu-rep --page source --csv --kernel-name regex:conv_igemm > $OUT/step_source_conv.csv 2> /dev/null
    rm -f $OUT/step.ncu-rep
    echo "report too large: kept csv pages only"
  fi
else
  echo "profile_step.py failed without ncu:"; tail -20 $OUT/profile_step.log
fi
