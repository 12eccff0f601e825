#!/bin/bash
# round-2 GPU call Z2 (1 GPU, short): register budget of the EPA kernel in the throughput regime (CLOSEST): 5 / 6 CTAs per SM (96 / 80 registers)
for V in 5 6; do
  T4C_LIB=$PWD/tesseract_b200/csrc/variants/libt4c_epa$V.so timeout 60 python bench.py --steps 2 --warmup 2 --no-secondary --no-cpu-baseline --mode closest > gpurun_out/bench_z2_epa$V.json 2> gpurun_out/bench_z2_epa$V.err; echo "epa $V exit $?"
done
