#!/bin/bash
# round-2 GPU call O (1 GPU): two-stage cull on / off, tiny last-round launch merged into the deferred pass
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py tests/test_gpu_fullsize.py -m gpu -q -x > gpurun_out/pytest_gpu_o.log 2>&1; echo "pytest exit $?"
tail -2 gpurun_out/pytest_gpu_o.log
for T in 1 0; do
  for M in first first_flags closest; do
    ST=10; [ $M = closest ] && ST=3
    T4C_TWO_STAGE_CULL=$T timeout 400 python bench.py --steps $ST --warmup 3 --no-secondary --no-cpu-baseline --mode $M > gpurun_out/bench_o_${M}_ts$T.json 2> gpurun_out/bench_o_${M}_ts$T.err; echo "$M two-stage=$T exit $?"
  done
done
T4C_TWO_STAGE_CULL=1 timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --secondary config2_cast_fine > gpurun_out/bench_o_cast_ts1.json 2> gpurun_out/bench_o_cast_ts1.err; echo "cast exit $?"
