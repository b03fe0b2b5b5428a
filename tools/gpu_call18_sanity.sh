#!/bin/bash
# last sanity of the final tree: parity tests of the nodal path, curved RHS / stepper / periodic maps, the no-alloc entry point (collected last), smoke
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_parity.py tests/test_curved_rhs.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | grep -v "^$" | tail -6 > gpurun_out/r02_last_sanity.log; tail -2 gpurun_out/r02_last_sanity.log
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1 | tee -a gpurun_out/r02_last_sanity.log
