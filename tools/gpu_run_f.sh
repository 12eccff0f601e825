#!/bin/bash
# round-2 GPU call F (8 GPUs): multi-GPU tests, scaling curve N = 1, 2, 4, 8 with the configs[4] entries
timeout 1500 python -m pytest tests/test_gpu_multi.py tests/test_cpp_managers.py tests/test_gpu_parity.py tests/test_gpu_golden.py tests/test_jacobian.py tests/test_scene_edit.py -m gpu -q -s > gpurun_out/pytest_gpu_f.log 2>&1; echo "pytest exit $?"
tail -4 gpurun_out/pytest_gpu_f.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err; echo "n1 exit $?"
for N in 2 4 8; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N \
    bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err; echo "n$N exit $?"
  tail -n 2 gpurun_out/scale_n$N.err
done
