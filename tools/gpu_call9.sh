#!/bin/bash
# round 2, call 9: per-node wall normals in the fused kernels -- parity, no perf regression; reference arm on the 16 M mesh
mkdir -p gpurun_out
rm -f gpurun_out/parity_fullsize.json
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_periodic_maps.py tests/test_coastal_workflow.py tests/test_python_reference_golden.py "tests/test_fullsize_parity.py::test_full_mesh_rhs_against_oracle_and_arbiter" -m gpu -q -k "not config5 and not config4" 2>&1 | grep -v "^$" | tail -8 > gpurun_out/r02_call9_pytest.log
tail -3 gpurun_out/r02_call9_pytest.log
timeout 400 python bench.py > gpurun_out/r02_bench_cfg2.json 2> gpurun_out/r02_bench_cfg2.err
timeout 300 python bench.py --mode full --no-cpu-baseline --no-e2e --no-parity --steps 5 > gpurun_out/r02_bench_full_quick.json 2>/dev/null
timeout 800 python bench.py --impl reference --gpus 8 --steps 3 --warmup 1 > gpurun_out/r02_bench_reference_n8.json 2> gpurun_out/r02_bench_reference_n8.err; echo "ref n8 rc=$?"
for f in cfg2 full_quick reference_n8; do python -c "
import json,sys
d=json.loads(open('gpurun_out/r02_bench_$f.json').read().strip().split('\n')[-1])
print('$f', round(d['value']/1e9,3), 'G', d.get('roofline',{}).get('frac'), 'ms_per_step', round(d['ms_per_step'],1), (d.get('cpu_baseline') or {}).get('sample','')[:200], d.get('max_rel_err_rhs'), (d.get('parity') or {}).get('geometry'))
" ; done
tail -3 gpurun_out/r02_bench_reference_n8.err
