#!/bin/bash
# Runs the GPU parity tests in separate processes per kernel group, so that a faulting
# kernel (sticky CUDA error) cannot mask the results of the other groups.
# Usage (on the GPU box): bash tests/run_gpu_groups.sh > gpurun_out/groups.log 2>&1
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv
for k in layout generic "tcgen05_3x3" tcx many_tiles plain_bf16 fp32_accurate dgrad wgrad maxpool gram test_mse instance tv flow_warp; do
  echo "=== group $k"
  timeout 600 python -m pytest tests/test_kernels_gpu.py -q -m gpu -k "$k" -p no:cacheprovider --tb=short 2>&1 | tail -40
done
echo "=== gatys"
timeout 900 python -m pytest tests/test_gatys_gpu.py -q -m gpu -p no:cacheprovider --tb=short 2>&1 | tail -60
echo "=== transform"
timeout 900 python -m pytest tests/test_transform_gpu.py -q -m gpu -p no:cacheprovider --tb=short 2>&1 | tail -80
echo "=== optim"
timeout 600 python -m pytest tests/test_optim_gpu.py -q -m gpu -p no:cacheprovider --tb=short 2>&1 | tail -40
echo "=== cli"
timeout 600 python -m pytest tests/test_cli_gpu.py -q -m gpu -p no:cacheprovider --tb=short 2>&1 | tail -40
