#!/bin/bash
# round-2 GPU call L (1 GPU): adaptive round slices (T4C_ROUND_MERGE) x round order (T4C_ROUND_ORDER), FIRST with records and flags-only
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/pytest_gpu_l.log 2>&1; echo "pytest exit $?"
tail -2 gpurun_out/pytest_gpu_l.log
for O in 1 0; do
  for MG in 0 35 70; do
    for M in first first_flags; do
      T4C_ROUND_ORDER=$O T4C_ROUND_MERGE=$MG timeout 300 python bench.py --steps 10 --warmup 3 --no-secondary --no-cpu-baseline --mode $M > gpurun_out/bench_l_${M}_o${O}_m${MG}.json 2> gpurun_out/bench_l_${M}_o${O}_m${MG}.err; echo "$M order=$O merge=$MG exit $?"
    done
  done
done
