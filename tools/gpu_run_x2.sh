#!/bin/bash
# round-2 GPU call X2 (1 GPU, short): last sanity of the final tree: smoke + one bench line
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_x2.log 2>&1; echo "smoke exit $?"
timeout 60 python bench.py --steps 5 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/bench_x2.json 2> gpurun_out/bench_x2.err; echo "bench exit $?"
