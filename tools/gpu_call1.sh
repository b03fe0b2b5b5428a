#!/bin/bash
# round 2, first GPU call: full GPU test suite (with the whole-mesh parity tables), the default bench line, configs 3/4 at their real sizes
mkdir -p gpurun_out
(nproc; free -g | head -2; nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv) > gpurun_out/r02_box.txt 2>&1
rm -f gpurun_out/parity_fullsize.json
python -m pytest tests -m gpu -q --durations=15 -s 2>&1 | grep -v "^$" | tail -80 > gpurun_out/r02_pytest_gpu.log
python bench.py > gpurun_out/r02_bench_cfg2.json 2> gpurun_out/r02_bench_cfg2.err
python bench.py --workload cfg3 --steps 3 > gpurun_out/r02_bench_cfg3.json 2> gpurun_out/r02_bench_cfg3.err
python bench.py --workload cfg4 --steps 3 > gpurun_out/r02_bench_cfg4.json 2> gpurun_out/r02_bench_cfg4.err
tail -5 gpurun_out/r02_pytest_gpu.log; cut -c1-300 gpurun_out/r02_bench_cfg2.json
