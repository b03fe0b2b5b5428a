#!/bin/bash
# Compiles the __host__ __device__ element functions of blitzdg_b200/csrc/sw2d_curved.cu into a small HOST test library
# (build/libcurved_host_emulation.so, test hook bdg_curved_host_emulation) for tests/test_curved_rhs.py.  Test
# infrastructure: never linked into libblitzdg_b200.so.
set -e
here=$(cd "$(dirname "$0")/.." && pwd)
mkdir -p $here/build
nccl=$(python -c "import nvidia.nccl; print(list(nvidia.nccl.__path__)[0])")
NVCC=$(command -v nvcc || echo /usr/local/cuda/bin/nvcc)
$NVCC -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -Xcompiler -fPIC -I$nccl/include -DBDG_CURVED_HOST_EMULATION \
  -shared -o $here/build/libcurved_host_emulation.so $here/blitzdg_b200/csrc/sw2d_curved.cu
