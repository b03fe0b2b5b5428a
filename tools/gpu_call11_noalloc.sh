#!/bin/bash
# bdg_sw2d_timestep_noalloc: alone in a fresh process, then in one process after the parity tests
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | grep -v "^$" | tail -15 > gpurun_out/r02_no_alloc_alone.log
tail -3 gpurun_out/r02_no_alloc_alone.log
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_python_reference_golden.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | grep -v "^$" | tail -15 > gpurun_out/r02_no_alloc_after_others.log
tail -3 gpurun_out/r02_no_alloc_after_others.log
