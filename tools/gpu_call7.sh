#!/bin/bash
# round 2, call 7: full-physics kernel with per-tile operator reload
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_coastal_workflow.py tests/test_periodic_maps.py -m gpu -q 2>&1 | tail -5 > gpurun_out/r02_call7_pytest.log
tail -3 gpurun_out/r02_call7_pytest.log
python bench.py --mode full --no-cpu-baseline --steps 5 --no-parity > gpurun_out/r02_bench_full_lserk4.json 2> gpurun_out/r02_bench_full_lserk4.err
python bench.py --mode full --stepper ssprk2 --no-cpu-baseline --steps 5 --no-parity > gpurun_out/r02_bench_coastal_ssprk2.json 2> gpurun_out/r02_bench_coastal_ssprk2.err
for f in full_lserk4 coastal_ssprk2; do python -c "
import json,sys
d=json.loads(open('gpurun_out/r02_bench_$f.json').read().strip().split('\n')[-1])
print('$f', round(d['value']/1e9,1), 'frac', round(d['roofline']['frac'],3), 'e2e', d['e2e'] and round(d['e2e']['value']/1e9,1), d['clocks']['sm_mhz'])
" ; done
ncu --set full --clock-control none --import-source on -k regex:k_fused_persistent -s 20 -c 2 -o gpurun_out/r02_fused_full_v2 -f python bench.py --mode full --steps 1 --warmup 1 --reps 1 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/ncu_f.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r02_launches_full_lserk4.csv python bench.py --mode full --steps 1 --warmup 1 --reps 1 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/ncu_l.log 2>&1
