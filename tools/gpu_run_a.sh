#!/bin/bash
# round-2 GPU call A: GPU test tier (all tests, no -x), the default bench line, the ncu launch list
export T4_PARITY_REPORT=gpurun_out/parity_reports.json T4_SECOND_OPINION_REPORT=gpurun_out/second_opinion.json
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv > gpurun_out/gpu.txt
lscpu | head -20 > gpurun_out/cpu.txt
python -m pytest tests -m gpu -q -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"
tail -5 gpurun_out/pytest_gpu.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_first.json 2> gpurun_out/bench_first.err; echo "bench exit $?"
tail -3 gpurun_out/bench_first.err
python bench.py --steps 2 --warmup 1 --no-secondary --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-secondary --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
echo "ncu exit $?"
