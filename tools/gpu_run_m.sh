#!/bin/bash
# round-2 GPU call M (1 GPU): whole GPU tier + default bench line of the current build; EPA block / type-sort hooks; ncu launch list
export T4_PARITY_REPORT=gpurun_out/parity_reports_m.json T4_SECOND_OPINION_REPORT=gpurun_out/second_opinion_m.json
timeout 1800 python -m pytest tests -m gpu -x -q -s > gpurun_out/pytest_gpu_m.log 2>&1; echo "pytest exit $?"
tail -3 gpurun_out/pytest_gpu_m.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_m.json 2> gpurun_out/bench_m.err; echo "bench exit $?"
T4C_EPA_BLOCK=128 timeout 300 python bench.py --steps 10 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/bench_m_epa128.json 2> gpurun_out/bench_m_epa128.err; echo "epa128 exit $?"
T4C_EPA_BLOCK=64 timeout 300 python bench.py --steps 10 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/bench_m_epa64.json 2> gpurun_out/bench_m_epa64.err; echo "epa64 exit $?"
T4C_SORT_MIN=1000000000 timeout 300 python bench.py --steps 10 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/bench_m_nosort.json 2> gpurun_out/bench_m_nosort.err; echo "nosort exit $?"
T4C_SORT_MIN=65536 timeout 300 python bench.py --steps 10 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/bench_m_sort64k.json 2> gpurun_out/bench_m_sort64k.err; echo "sort64k exit $?"
python bench.py --steps 1 --warmup 1 --no-secondary --no-cpu-baseline > gpurun_out/plain_m.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_m.csv \
    python bench.py --steps 1 --warmup 1 --no-secondary --no-cpu-baseline > gpurun_out/ncu_launch_m.log 2>&1
echo "ncu list exit $?"
