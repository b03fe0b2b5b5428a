#!/bin/bash
# round 2, final validation: the driver's own sequence (GPU tests, smoke, bench, reference arm)
mkdir -p gpurun_out
rm -f gpurun_out/parity_fullsize.json
timeout 1500 python -m pytest tests -m gpu -q --durations=8 2>&1 | grep -v "^$" | tail -30 > gpurun_out/r02_final_pytest_gpu.log
tail -4 gpurun_out/r02_final_pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_final_smoke.log 2>&1; tail -1 gpurun_out/r02_final_smoke.log
timeout 400 python bench.py > gpurun_out/r02_bench_cfg2.json 2> gpurun_out/r02_bench_cfg2.err
timeout 400 python bench.py --impl reference > gpurun_out/r02_bench_reference_n1.json 2> gpurun_out/r02_bench_reference_n1.err
/usr/bin/time -v timeout 800 python bench.py --impl reference --gpus 8 --steps 3 --warmup 1 > gpurun_out/r02_bench_reference_n8.json 2> gpurun_out/r02_bench_reference_n8.err
grep -E "Maximum resident|Elapsed" gpurun_out/r02_bench_reference_n8.err
for f in cfg2 reference_n1 reference_n8; do python -c "
import json,sys
d=json.loads(open('gpurun_out/r02_bench_$f.json').read().strip().split('\n')[-1])
print('$f', round(d['value']/1e9,3), 'G', d.get('roofline',{}).get('frac'), 'ms_per_step', round(d['ms_per_step'],1), (d.get('cpu_baseline') or {}).get('sample','')[:160], d.get('max_rel_err_rhs'), d.get('max_rel_err_1000'))
" ; done
