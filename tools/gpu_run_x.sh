#!/bin/bash
# round-2 GPU call X (1 GPU, short): sanity of the last build (t4c_build_info added): smoke, golden + no-alloc tests, one bench line
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_x.log 2>&1; echo "smoke exit $?"
timeout 150 python -m pytest tests/test_gpu_golden.py tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_gpu_x.log 2>&1; echo "pytest exit $?"
tail -1 gpurun_out/pytest_gpu_x.log
timeout 100 python bench.py --steps 5 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/bench_x.json 2> gpurun_out/bench_x.err; echo "bench exit $?"
