#!/bin/bash
# residual in C-fragment order (-DBDG_RES_FRAG, libblitzdg_b200_rfrag.so) against the default build: parity, then alternating benches
mkdir -p gpurun_out
V=$PWD/blitzdg_b200/csrc/libblitzdg_b200_rfrag.so
T="tests/test_gpu_parity.py tests/test_periodic_maps.py"
timeout 400 python -m pytest $T -m gpu -x -q 2>&1 | grep -v "^$" | tail -6 > gpurun_out/r02_rfrag_pytest_base.log; tail -2 gpurun_out/r02_rfrag_pytest_base.log
BDG_B200_LIB=$V timeout 500 python -m pytest $T "tests/test_fullsize_parity.py::test_full_mesh_rhs_against_oracle_and_arbiter" "tests/test_fullsize_parity.py::test_full_size_properties" -k "not config3 and not config4 and not config5" -m gpu -x -q 2>&1 | grep -v "^$" | tail -12 > gpurun_out/r02_rfrag_pytest_variant.log; tail -3 gpurun_out/r02_rfrag_pytest_variant.log
Q="--steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-parity"
for rep in 1 2; do
  for name in base rfrag; do
    if [ $name = base ]; then lib=""; else lib=$V; fi
    BDG_B200_LIB=$lib timeout 150 python bench.py $Q 2> gpurun_out/err_$name.log | tail -1 > gpurun_out/r02_rfrag_bench_${name}_$rep.json
    python -c "
import json; d=json.loads(open('gpurun_out/r02_rfrag_bench_${name}_$rep.json').read())
print('$name $rep', round(d['value']/1e9,2), 'G frac', round(d['roofline']['frac'],4), 'best', round(d['repetitions']['best']/1e9,2), d['clocks']['sm_mhz'], d['clocks']['power_w'], d['clocks']['reasons'], 'eta', d.get('eta_max_after'))"
  done
done
BDG_B200_LIB=$V timeout 400 python bench.py --no-cpu-baseline > gpurun_out/r02_rfrag_bench_full_line.json 2> gpurun_out/err_rfrag_full.log
python -c "
import json; d=json.loads(open('gpurun_out/r02_rfrag_bench_full_line.json').read().strip().split('\n')[-1])
print('rfrag full line', round(d['value']/1e9,2), 'G', round(d['roofline']['frac'],4), 'e2e', round(d['e2e']['value']/1e9,2), 'err_rhs', d.get('max_rel_err_rhs'), 'err_1000', d.get('max_rel_err_1000'))"
