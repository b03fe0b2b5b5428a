#!/bin/bash
# round 2, call 4: two compute warps per element group at N >= 6
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_periodic_maps.py tests/test_coastal_workflow.py -m gpu -q 2>&1 | tail -30 > gpurun_out/r02_call4_pytest.log
tail -3 gpurun_out/r02_call4_pytest.log
python bench.py --workload cfg3 --steps 3 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/r02_bench_cfg3_fused_wh2.json 2> gpurun_out/r02_bench_cfg3_fused_wh2.err
python bench.py --order 6 --nx 1000 --ny 500 --steps 3 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/r02_bench_tri6_fused_wh2.json 2>/dev/null
python bench.py --order 7 --nx 1000 --ny 500 --steps 3 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/r02_bench_tri7_fused_wh2.json 2>/dev/null
python bench.py --order 7 --nx 1000 --ny 500 --steps 3 --variant 1 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/r02_bench_tri7_v1.json 2>/dev/null
for f in cfg3_fused_wh2 tri6_fused_wh2 tri7_fused_wh2 tri7_v1; do python -c "
import json,sys
d=json.loads(open('gpurun_out/r02_bench_$f.json').read().strip().split('\n')[-1])
print('$f', round(d['value']/1e9,1), 'frac', round(d['roofline']['frac'],3), 'fp64', round(d['fp64_fraction'],3), 'var', d['config']['kernel_variant'], d['clocks']['sm_mhz'])
" ; done
ncu --set full --clock-control none --import-source on -k regex:k_fused_persistent -s 10 -c 1 -o gpurun_out/r02_fused_n8_wh2 -f python bench.py --workload cfg3 --steps 1 --warmup 1 --reps 1 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/ncu_n8.log 2>&1
