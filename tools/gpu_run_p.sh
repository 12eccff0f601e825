#!/bin/bash
# round-2 GPU call P (1 GPU): the workloads of ranks 0..7 of the scaling run (seed 1234 + rank) one after the other on one GPU:
# the step time of the slowest is what bounds the N-GPU step
for R in 0 1 2 3 4 5 6 7; do
  T4_BENCH_SEED_OFFSET=$R timeout 300 python bench.py --steps 10 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/bench_p_seed$R.json 2> gpurun_out/bench_p_seed$R.err; echo "seed offset $R exit $?"
done
