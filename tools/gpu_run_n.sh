#!/bin/bash
# round-2 GPU call N (1 GPU): ncu --set full of one step of the current build (all five kernels), source view of cull + GJK
python bench.py --steps 1 --warmup 1 --no-secondary --no-cpu-baseline > gpurun_out/plain_n.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:"gjkKernelR128|epaKernelR128|cullKernel|broadStaticKernel|fkKernelT|liveKeyKernel" \
    -s 25 -c 25 -f -o gpurun_out/prof_n python bench.py --steps 1 --warmup 1 --no-secondary --no-cpu-baseline > gpurun_out/ncu_full_n.log 2>&1
echo "ncu full exit $?"
ncu -i gpurun_out/prof_n.ncu-rep --page raw --csv > gpurun_out/n_full_raw.csv 2> gpurun_out/ncu_export_n.err
ncu -i gpurun_out/prof_n.ncu-rep --page source --csv --print-source sass -k regex:"cullKernel" > gpurun_out/n_cull_source_sass.csv 2>> gpurun_out/ncu_export_n.err
ls -la gpurun_out/
rm -f gpurun_out/prof_n.ncu-rep
