#!/bin/bash
# round-2 GPU call Z (1 GPU, short): resident CTAs per SM of the persistent EPA grid in the throughput regime (CLOSEST: 2.6 M tasks)
for V in 4 3 2; do
  T4C_EPA_CTAS=$V timeout 60 python bench.py --steps 2 --warmup 2 --no-secondary --no-cpu-baseline --mode closest > gpurun_out/bench_z_ctas$V.json 2> gpurun_out/bench_z_ctas$V.err; echo "ctas $V exit $?"
done
