#!/bin/bash
# round-2 GPU call Q (1 GPU): small EPA launches with their polytopes in a global pool (T4C_EPA_POOL) on / off, over the
# workloads of ranks 0, 3, 5 of the scaling run (their EPA tails were 1.48 / 1.88 / 2.03 ms in r3p)
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "first or config2 or flags or golden" > gpurun_out/pytest_gpu_q.log 2>&1; echo "pytest exit $?"
tail -2 gpurun_out/pytest_gpu_q.log
for R in 0 3 5; do
  for E in 1 0; do
    T4C_EPA_POOL=$E T4_BENCH_SEED_OFFSET=$R timeout 300 python bench.py --steps 10 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/bench_q_seed${R}_pool$E.json 2> gpurun_out/bench_q_seed${R}_pool$E.err; echo "seed $R pool $E exit $?"
  done
done
