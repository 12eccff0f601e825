#!/bin/bash
# round-2 GPU call V (1 GPU): curved penetrating pairs in the first deferred pass (T4C_CURVED_FIRST) on / off
timeout 400 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "first or flags or config2" > gpurun_out/pytest_gpu_v.log 2>&1; echo "pytest exit $?"
tail -2 gpurun_out/pytest_gpu_v.log
for R in 0 3 5; do
  for E in 1 0; do
    T4C_CURVED_FIRST=$E T4_BENCH_SEED_OFFSET=$R timeout 200 python bench.py --steps 10 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/bench_v_seed${R}_cf$E.json 2> gpurun_out/bench_v_seed${R}_cf$E.err; echo "seed $R curved-first $E exit $?"
  done
done
