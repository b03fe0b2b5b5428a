#!/bin/bash
# round 2, final-tree validation: the driver's own sequence (GPU tests, smoke, default bench), then the ncu launch list
# and one --set full capture of the headline kernel.  Each ncu pass only after its command has exited 0 without ncu.
mkdir -p gpurun_out
rm -f gpurun_out/parity_fullsize.json
if [ -f tests/test_zz_no_alloc.py ]; then  # alone, in a fresh process, before anything else has touched the GPU
  timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | grep -v "^$" | tail -6 > gpurun_out/r02_no_alloc_alone.log
  tail -2 gpurun_out/r02_no_alloc_alone.log
fi
timeout 1300 python -m pytest tests -m gpu -x -q --durations=8 2>&1 | grep -v "^$" | tail -30 > gpurun_out/r02_final_pytest_gpu.log
tail -4 gpurun_out/r02_final_pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_final_smoke.log 2>&1; tail -1 gpurun_out/r02_final_smoke.log
timeout 400 python bench.py > gpurun_out/r02_bench_cfg2.json 2> gpurun_out/r02_bench_cfg2.err; rc=$?
python -c "
import json
d=json.loads(open('gpurun_out/r02_bench_cfg2.json').read().strip().split('\n')[-1])
print('cfg2', round(d['value']/1e9,3), 'G', d['roofline']['frac'], 'e2e', round(d['e2e']['value']/1e9,2), d.get('max_rel_err_rhs'), (d.get('parity') or {}).get('geometry'))
"
if [ $rc -eq 0 ]; then
  Q="--steps 2 --warmup 1 --reps 1 --no-cpu-baseline --no-e2e --no-parity"
  timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_cfg2.csv \
    python bench.py $Q > gpurun_out/ncu_list.log 2>&1
  timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_fused_persistent -s 20 -c 2 -f -o gpurun_out/r02_fused_cfg2 \
    python bench.py $Q > gpurun_out/ncu_full.log 2>&1
  ls -la gpurun_out/*.ncu-rep
fi
