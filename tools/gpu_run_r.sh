#!/bin/bash
# round-2 GPU call R (8 GPUs): multi-GPU tests on 4 devices; scaling curve N = 1, 2, 4, 8 of the final build (configs[4] entries at N = 8)
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -q -s > gpurun_out/pytest_gpu_r.log 2>&1; echo "pytest exit $?"
tail -3 gpurun_out/pytest_gpu_r.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/scale_r_n1.json 2> gpurun_out/scale_r_n1.err; echo "n1 exit $?"
for N in 8 4 2; do
  EXTRA="--no-secondary"; [ $N = 8 ] && EXTRA=""
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2971$N \
    bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline $EXTRA > gpurun_out/scale_r_n$N.json 2> gpurun_out/scale_r_n$N.err; echo "n$N exit $?"
  tail -n 1 gpurun_out/scale_r_n$N.err
done
T4C_PEER_PUSH=0 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29720 \
    bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu-baseline --no-secondary > gpurun_out/scale_r_n8_nopush.json 2> gpurun_out/scale_r_n8_nopush.err; echo "n8 nopush exit $?"
