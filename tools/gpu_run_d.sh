#!/bin/bash
# round-2 GPU call D (2 GPUs): full GPU test tier, bench N=1 (full line) and N=2
export T4_PARITY_REPORT=gpurun_out/parity_reports.json T4_SECOND_OPINION_REPORT=gpurun_out/second_opinion.json
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_d.log 2>&1; echo "pytest exit $?"
tail -6 gpurun_out/pytest_gpu_d.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_d_n1.json 2> gpurun_out/bench_d_n1.err; echo "bench n1 exit $?"
tail -n 3 gpurun_out/bench_d_n1.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_d_n2.json 2> gpurun_out/bench_d_n2.err; echo "bench n2 exit $?"
tail -n 5 gpurun_out/bench_d_n2.err
