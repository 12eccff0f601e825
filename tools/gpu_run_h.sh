#!/bin/bash
# round-2 GPU call H (1 GPU): deferred tiers share one EPA launch; source-level capture of the largest GJK launch
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/pytest_gpu_h.log 2>&1; echo "pytest exit $?"
tail -3 gpurun_out/pytest_gpu_h.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/bench_h.json 2> gpurun_out/bench_h.err; echo "bench exit $?"
python bench.py --steps 1 --warmup 1 --no-secondary --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"gjkKernelR128" -s 12 -c 2 -f -o gpurun_out/prof_h \
    python bench.py --steps 1 --warmup 1 --no-secondary --no-cpu-baseline > gpurun_out/ncu_h.log 2>&1
echo "ncu exit $?"
ncu -i gpurun_out/prof_h.ncu-rep --page raw --csv > gpurun_out/h_gjk_raw.csv 2> gpurun_out/ncu_export.err
ncu -i gpurun_out/prof_h.ncu-rep --page source --csv --print-source sass > gpurun_out/h_gjk_source_sass.csv 2>> gpurun_out/ncu_export.err
ncu -i gpurun_out/prof_h.ncu-rep --page source --csv --print-source cuda > gpurun_out/h_gjk_source_cuda.csv 2>> gpurun_out/ncu_export.err
ls -la gpurun_out/
rm -f gpurun_out/prof_h.ncu-rep
