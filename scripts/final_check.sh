#!/bin/bash
# Last single-GPU call of the round (about 2.5 minutes of box time): GPU tests, one bench line, one ncu capture.
OUT=gpurun_out/final; mkdir -p $OUT
timeout 45 python -m pytest tests -q -x -m "gpu and not multigpu" > $OUT/tests_1gpu.log 2>&1; tail -2 $OUT/tests_1gpu.log
timeout 60 python bench.py --steps 100 --warmup 10 --no-strong > $OUT/bench_1gpu_native.json 2> $OUT/bench_1gpu_native.err
tail -c 900 $OUT/bench_1gpu_native.json
if timeout 30 python scripts/profile_step.py --select > $OUT/profile_step.log 2>&1; then
  timeout 70 ncu --set full --clock-control none --import-source on --profile-from-start off -f -o $OUT/step \
      python scripts/profile_step.py --select > $OUT/ncu.log 2>&1
  tail -2 $OUT/ncu.log
  ls -la $OUT
fi
