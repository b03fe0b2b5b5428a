#!/bin/bash
# over-integrated RHS on curved triangles through the C ABI against the reference fixture and the oracle
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_curved_rhs.py -m gpu -q 2>&1 | grep -v "^$" | tail -25 > gpurun_out/r02_curved_pytest.log
tail -12 gpurun_out/r02_curved_pytest.log
