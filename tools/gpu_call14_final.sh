#!/bin/bash
# round 2, final tree (fragment-ordered residual, no-alloc entry point, curved RHS): the driver's sequence, ncu evidence of the
# headline kernel, then an A/B of three build variants of that kernel on the same box.
mkdir -p gpurun_out
rm -f gpurun_out/parity_fullsize.json
timeout 200 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | grep -v "^$" | tail -4 > gpurun_out/r02_no_alloc_alone.log; tail -1 gpurun_out/r02_no_alloc_alone.log
# configs 3 and 4 at full size (N = 8 triangles, quads: kernels unchanged since their validation in r02_final_pytest_gpu.log) are left
# to the driver's own run: 4 minutes of CPU arbiter time on the GPU box
timeout 900 python -m pytest tests -m gpu -x -q --durations=6 -k "not config3 and not config4" 2>&1 | grep -v "^$" | tail -16 > gpurun_out/r02_final2_pytest_gpu.log
tail -3 gpurun_out/r02_final2_pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_final_smoke.log 2>&1; tail -1 gpurun_out/r02_final_smoke.log
timeout 400 python bench.py > gpurun_out/r02_bench_cfg2.json 2> gpurun_out/r02_bench_cfg2.err; rc=$?
show() { python -c "
import json,sys
d=json.loads(open('$1').read().strip().split('\n')[-1])
print('$2', round(d['value']/1e9,2), 'G frac', round(d['roofline']['frac'],4), 'best', round(d['repetitions']['best']/1e9,2), d['clocks']['sm_mhz'], d['clocks'].get('power_w'), d['clocks']['reasons'], 'e2e', round((d.get('e2e') or {}).get('value',0)/1e9,2), d.get('max_rel_err_rhs'), d.get('max_rel_err_1000'))"; }
show gpurun_out/r02_bench_cfg2.json final
Q="--steps 2 --warmup 1 --reps 1 --no-cpu-baseline --no-e2e --no-parity"
if [ $rc -eq 0 ]; then
  timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_cfg2.csv \
    python bench.py $Q > gpurun_out/ncu_list.log 2>&1
  timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_fused_persistent -s 20 -c 2 -f -o gpurun_out/r02_fused_cfg2 \
    python bench.py $Q > gpurun_out/ncu_full.log 2>&1
  ls -la gpurun_out/*.ncu-rep
fi
Q="--steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-parity"
for name in base padio nopf staged base padio; do
  if [ $name = base ]; then lib=""; else lib=$PWD/blitzdg_b200/csrc/libblitzdg_b200_$name.so; fi
  BDG_B200_LIB=$lib timeout 150 python bench.py $Q 2> gpurun_out/err_$name.log | tail -1 > gpurun_out/r02_ab2_$name.json
  show gpurun_out/r02_ab2_$name.json $name
  cat gpurun_out/r02_ab2_$name.json >> gpurun_out/r02_ab2_all.jsonl
done
