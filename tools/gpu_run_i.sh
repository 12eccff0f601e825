#!/bin/bash
# round-2 GPU call I (2 GPUs): gather slab + progressive peer-memory push (threads and processes), N = 2 bench with / without
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q -s -x > gpurun_out/pytest_gpu_i.log 2>&1; echo "pytest exit $?"
tail -5 gpurun_out/pytest_gpu_i.log
for PUSH in 1 0; do
  T4C_PEER_PUSH=$PUSH timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2961$PUSH \
    bench.py --gpus 2 --steps 10 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/bench_i_n2_push$PUSH.json 2> gpurun_out/bench_i_n2_push$PUSH.err; echo "n2 push=$PUSH exit $?"
  tail -n 2 gpurun_out/bench_i_n2_push$PUSH.err
done
