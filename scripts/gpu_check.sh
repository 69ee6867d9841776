#!/usr/bin/env bash
# Run on the GPU box via gpurun: GPU tests (not stopping at the first failure) with logs kept.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout ${1:-900} python -m pytest tests -m gpu -q -x --timeout=300 ${2:-} > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -40 gpurun_out/pytest_gpu.log
