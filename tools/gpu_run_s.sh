#!/bin/bash
# round-2 GPU call S (1 GPU): EPA tasks of the first deferred pass start early on their own stream (T4C_EPA_EARLY) on / off
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_cpp_managers.py -m gpu -q -x > gpurun_out/pytest_gpu_s.log 2>&1; echo "pytest exit $?"
tail -2 gpurun_out/pytest_gpu_s.log
for R in 0 3 5; do
  for E in 1 0; do
    T4C_EPA_EARLY=$E T4_BENCH_SEED_OFFSET=$R timeout 300 python bench.py --steps 10 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/bench_s_seed${R}_early$E.json 2> gpurun_out/bench_s_seed${R}_early$E.err; echo "seed $R early $E exit $?"
  done
done
