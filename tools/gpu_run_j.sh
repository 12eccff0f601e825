#!/bin/bash
# round-2 GPU call J (1 GPU): t4c_fk_noalloc test alone and after the whole GPU tier; default bench line; variant builds
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_gpu_j_noalloc_alone.log 2>&1; echo "noalloc alone exit $?"
tail -2 gpurun_out/pytest_gpu_j_noalloc_alone.log
export T4_PARITY_REPORT=gpurun_out/parity_reports_j.json T4_SECOND_OPINION_REPORT=gpurun_out/second_opinion_j.json
timeout 1800 python -m pytest tests -m gpu -x -q -s > gpurun_out/pytest_gpu_j.log 2>&1; echo "pytest exit $?"
tail -3 gpurun_out/pytest_gpu_j.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_j.json 2> gpurun_out/bench_j.err; echo "bench exit $?"
tail -c 300 gpurun_out/bench_j.err
for V in base tetra scan4 both; do
  if [ $V = base ]; then unset T4C_LIB; else export T4C_LIB=$PWD/tesseract_b200/csrc/variants/libt4c_$V.so; fi
  timeout 400 python bench.py --steps 10 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/bench_j_$V.json 2> gpurun_out/bench_j_$V.err; echo "variant $V exit $?"
  timeout 400 python bench.py --steps 3 --warmup 2 --no-secondary --no-cpu-baseline --mode closest > gpurun_out/bench_j_${V}_closest.json 2> gpurun_out/bench_j_${V}_closest.err; echo "variant $V closest exit $?"
done
