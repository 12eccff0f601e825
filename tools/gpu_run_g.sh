#!/bin/bash
# round-2 GPU call G (1 GPU): certified tier 0 (cull keeps iterating until the cores are known to be apart)
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py tests/test_gpu_fullsize.py -m gpu -q -s -x > gpurun_out/pytest_gpu_g.log 2>&1; echo "pytest exit $?"
tail -4 gpurun_out/pytest_gpu_g.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_g.json 2> gpurun_out/bench_g.err; echo "bench exit $?"
tail -c 600 gpurun_out/bench_g.err
T4C_TRACE_PASSES=1 timeout 300 python bench.py --steps 1 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/bench_g_trace.json 2> gpurun_out/bench_g_trace.err; echo "trace exit $?"
