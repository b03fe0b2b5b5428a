#!/bin/bash
# round 2, final tree after the stage-form dispatch fix (LSERK4 -> fragment-ordered residual kernels, SSP-RK2 -> staged kernels)
mkdir -p gpurun_out
rm -f gpurun_out/parity_fullsize.json
# configs 3 and 4 at full size (N = 8 triangles, quads: kernels unchanged since r02_final_pytest_gpu.log) are left to the driver's run
timeout 1000 python -m pytest tests -m gpu -q --durations=6 -k "not config3 and not config4" 2>&1 | grep -v "^$" | tail -25 > gpurun_out/r02_final2_pytest_gpu.log
tail -4 gpurun_out/r02_final2_pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_final_smoke.log 2>&1; tail -1 gpurun_out/r02_final_smoke.log
timeout 400 python bench.py > gpurun_out/r02_bench_cfg2.json 2> gpurun_out/r02_bench_cfg2.err
python -c "
import json
d=json.loads(open('gpurun_out/r02_bench_cfg2.json').read().strip().split('\n')[-1])
print('final', round(d['value']/1e9,2), 'G frac', round(d['roofline']['frac'],4), 'best', round(d['repetitions']['best']/1e9,2), d['clocks']['sm_mhz'], d['clocks']['reasons'], 'e2e', round(d['e2e']['value']/1e9,2), d.get('max_rel_err_rhs'), d.get('max_rel_err_1000'))"
