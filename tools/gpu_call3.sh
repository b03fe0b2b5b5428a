#!/bin/bash
# round 2, call 3: fused kernel at N=6..8, tuned full-physics kernel
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_coastal_workflow.py tests/test_python_reference_golden.py tests/test_periodic_maps.py -m gpu -q 2>&1 | tail -30 > gpurun_out/r02_call3_pytest.log
tail -3 gpurun_out/r02_call3_pytest.log
python bench.py --workload cfg3 --steps 3 --no-cpu-baseline > gpurun_out/r02_bench_cfg3_fused.json 2> gpurun_out/r02_bench_cfg3_fused.err
python bench.py --order 6 --nx 1000 --ny 500 --steps 3 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/r02_bench_tri6_fused.json 2>/dev/null
python bench.py --order 6 --nx 1000 --ny 500 --steps 3 --variant 1 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/r02_bench_tri6_v1.json 2>/dev/null
python bench.py --order 5 --nx 1000 --ny 500 --steps 3 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/r02_bench_tri5_fused.json 2>/dev/null
python bench.py --mode full --no-cpu-baseline --steps 5 --no-e2e --no-parity > gpurun_out/r02_bench_full_lserk4.json 2> gpurun_out/r02_bench_full_lserk4.err
python bench.py --mode full --stepper ssprk2 --no-cpu-baseline --steps 5 --no-e2e --no-parity > gpurun_out/r02_bench_coastal_ssprk2.json 2> gpurun_out/r02_bench_coastal_ssprk2.err
python bench.py --no-cpu-baseline --no-e2e --no-parity > gpurun_out/r02_bench_cfg2_quick.json 2>/dev/null
for f in cfg3_fused tri6_fused tri6_v1 tri5_fused full_lserk4 coastal_ssprk2 cfg2_quick; do python -c "
import json,sys
d=json.loads(open('gpurun_out/r02_bench_$f.json').read().strip().split('\n')[-1])
print('$f', round(d['value']/1e9,1), 'frac', round(d['roofline']['frac'],3), 'fp64', round(d['fp64_fraction'],3), 'var', d['config']['kernel_variant'], d['clocks']['sm_mhz'])
" ; done
ncu --set full --clock-control none --import-source on -k regex:k_fused_persistent -s 10 -c 1 -o gpurun_out/r02_fused_n8 -f python bench.py --workload cfg3 --steps 1 --warmup 1 --reps 1 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/ncu_n8.log 2>&1
