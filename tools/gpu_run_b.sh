#!/bin/bash
# round-2 GPU call B: cooperative shared-memory EPA -- parity tier + headline bench (FIRST, CLOSEST)
export T4_PARITY_REPORT=gpurun_out/parity_reports_b.json
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py tests/test_cpp_managers.py -m gpu -q -x > gpurun_out/pytest_gpu_b.log 2>&1; echo "pytest exit $?"
tail -4 gpurun_out/pytest_gpu_b.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-secondary > gpurun_out/bench_b_first.json 2> gpurun_out/bench_b_first.err; echo "bench first exit $?"
timeout 600 python bench.py --steps 3 --warmup 2 --no-secondary --mode closest --no-cpu-baseline > gpurun_out/bench_b_closest.json 2> gpurun_out/bench_b_closest.err; echo "bench closest exit $?"
tail -2 gpurun_out/bench_b_first.err gpurun_out/bench_b_closest.err
