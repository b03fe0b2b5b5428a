#!/bin/bash
# warp-local phase 1 + per-warp `done` arrivals (-DBDG_WARP_LOCAL_PHASE1): parity with the variant library, then A/B on one box
mkdir -p gpurun_out
V=$PWD/blitzdg_b200/csrc/libblitzdg_b200_wlocal.so
BDG_B200_LIB=$V timeout 400 python -m pytest tests/test_gpu_parity.py tests/test_periodic_maps.py tests/test_coastal_workflow.py tests/test_python_reference_golden.py "tests/test_fullsize_parity.py::test_full_mesh_rhs_against_oracle_and_arbiter" "tests/test_fullsize_parity.py::test_full_size_properties" -k "not config3 and not config4 and not config5" -m gpu -q 2>&1 | grep -v "^$" | tail -8 > gpurun_out/r02_wlocal_pytest.log; tail -2 gpurun_out/r02_wlocal_pytest.log
Q="--steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-parity"
for name in wlocal base wlocal; do
  if [ $name = base ]; then lib=""; else lib=$V; fi
  BDG_B200_LIB=$lib timeout 150 python bench.py $Q 2> gpurun_out/err_$name.log | tail -1 > gpurun_out/tmp.json
  cat gpurun_out/tmp.json >> gpurun_out/r02_wlocal_ab.jsonl
  python -c "
import json; d=json.loads(open('gpurun_out/tmp.json').read())
print('$name', round(d['value']/1e9,2), 'G frac', round(d['roofline']['frac'],4), 'best', round(d['repetitions']['best']/1e9,2), d['clocks']['sm_mhz'], d['clocks'].get('power_w'), d['clocks']['reasons'], 'eta', d.get('eta_max_after'))"
done
