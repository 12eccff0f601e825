#!/bin/bash
# round-2 GPU call W (2 GPUs, short): multi-GPU tests and one N = 2 bench with the configs[4] entries (slab growth between modes)
timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -q -x > gpurun_out/pytest_gpu_w.log 2>&1; echo "pytest exit $?"
tail -2 gpurun_out/pytest_gpu_w.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29741 \
    bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu-baseline --secondary config4_dual_arm_first,config4_dual_arm_closest > gpurun_out/bench_w_n2.json 2> gpurun_out/bench_w_n2.err; echo "n2 exit $?"
tail -n 2 gpurun_out/bench_w_n2.err
