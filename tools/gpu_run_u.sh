#!/bin/bash
# round-2 GPU call U (8 GPUs, short): N = 8 and N = 1 of the final tree (EPA tasks start early: the rank-to-rank spread)
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29731 \
    bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu-baseline --no-secondary > gpurun_out/scale_u_n8.json 2> gpurun_out/scale_u_n8.err; echo "n8 exit $?"
timeout 100 python bench.py --steps 10 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/scale_u_n1.json 2> gpurun_out/scale_u_n1.err; echo "n1 exit $?"
