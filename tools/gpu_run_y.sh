#!/bin/bash
# round-2 GPU call Y (1 GPU, short): register budget of the exact GJK kernel: 3 / 4 / 5 / 6 CTAs of 128 threads per SM (168 / 128 / 96 / 80 registers)
for V in 3 5 6; do
  T4C_LIB=$PWD/tesseract_b200/csrc/variants/libt4c_gjk$V.so timeout 60 python bench.py --steps 5 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/bench_y_gjk$V.json 2> gpurun_out/bench_y_gjk$V.err; echo "gjk $V exit $?"
done
T4C_LIB=$PWD/tesseract_b200/csrc/variants/libt4c_gjk5.so timeout 60 python bench.py --steps 2 --warmup 2 --no-secondary --no-cpu-baseline --mode closest > gpurun_out/bench_y_gjk5_closest.json 2> gpurun_out/bench_y_gjk5_closest.err; echo "closest 5 exit $?"
