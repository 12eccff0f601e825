#!/bin/bash
# round-2 GPU call T (1 GPU): whole GPU tier + smoke + default bench line of the final tree; ncu launch list and full capture
export T4_PARITY_REPORT=gpurun_out/parity_reports_t.json T4_SECOND_OPINION_REPORT=gpurun_out/second_opinion_t.json
timeout 1800 python -m pytest tests -m gpu -x -q -s > gpurun_out/pytest_gpu_t.log 2>&1; echo "pytest exit $?"
tail -3 gpurun_out/pytest_gpu_t.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke_t.log 2>&1; echo "smoke exit $?"
timeout 900 python bench.py > gpurun_out/bench_t.json 2> gpurun_out/bench_t.err; echo "bench (default flags) exit $?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_t_reference.json 2> gpurun_out/bench_t_reference.err; echo "reference arm exit $?"
timeout 300 python bench.py --steps 10 --warmup 3 --no-secondary --no-cpu-baseline --mode first_flags > gpurun_out/bench_t_flags.json 2> gpurun_out/bench_t_flags.err; echo "flags exit $?"
python bench.py --steps 1 --warmup 1 --no-secondary --no-cpu-baseline > gpurun_out/plain_t.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_t.csv \
    python bench.py --steps 1 --warmup 1 --no-secondary --no-cpu-baseline > gpurun_out/ncu_launch_t.log 2>&1
echo "ncu list exit $?"
