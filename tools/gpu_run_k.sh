#!/bin/bash
# round-2 GPU call K (1 GPU): rounds walk a Morton-ordered list of the live states (T4C_LIVE_ORDER), on / off
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_cpp_managers.py -m gpu -q -x > gpurun_out/pytest_gpu_k.log 2>&1; echo "pytest exit $?"
tail -3 gpurun_out/pytest_gpu_k.log
for L in 1 0; do
  for M in first first_flags; do
    T4C_LIVE_ORDER=$L timeout 400 python bench.py --steps 10 --warmup 3 --no-secondary --no-cpu-baseline --mode $M > gpurun_out/bench_k_${M}_live$L.json 2> gpurun_out/bench_k_${M}_live$L.err; echo "$M live=$L exit $?"
  done
done
