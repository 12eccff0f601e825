#!/bin/bash
# round-2 GPU call E (1 GPU): full GPU test tier, the default bench line, ncu launch list + full capture of one step
export T4_PARITY_REPORT=gpurun_out/parity_reports.json T4_SECOND_OPINION_REPORT=gpurun_out/second_opinion.json
timeout 1800 python -m pytest tests -m gpu -q -s > gpurun_out/pytest_gpu_e.log 2>&1; echo "pytest exit $?"
tail -4 gpurun_out/pytest_gpu_e.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_e_n1.json 2> gpurun_out/bench_e_n1.err; echo "bench n1 exit $?"
tail -n 3 gpurun_out/bench_e_n1.err
timeout 300 python bench.py --steps 3 --warmup 2 --no-secondary --no-cpu-baseline --mode first_flags > gpurun_out/bench_e_flags.json 2> gpurun_out/bench_e_flags.err; echo "bench flags exit $?"
python bench.py --steps 1 --warmup 1 --no-secondary --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 1 --warmup 1 --no-secondary --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
echo "ncu list exit $?"
ncu --set full --clock-control none --import-source on -k regex:"gjkKernelR128|epaKernelR128|cullKernel|broadStaticKernel|fkKernelT|pruneKernel|broadActiveKernel" \
    -s 26 -c 30 -f -o gpurun_out/prof python bench.py --steps 1 --warmup 1 --no-secondary --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?"
# the report itself is too large to travel back (64 MiB limit on gpurun_out/): keep its raw page as CSV
ncu -i gpurun_out/prof.ncu-rep --page raw --csv > gpurun_out/full_raw.csv 2> gpurun_out/ncu_export.err
ls -la gpurun_out/prof.ncu-rep gpurun_out/full_raw.csv
rm -f gpurun_out/prof.ncu-rep
