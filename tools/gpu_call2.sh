#!/bin/bash
# round 2, call 2: fused full-physics kernel -- parity, memcheck, bench lines, ncu
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_coastal_workflow.py tests/test_periodic_maps.py tests/test_python_reference_golden.py tests/test_output_path.py -m gpu -q 2>&1 | tail -40 > gpurun_out/r02_call2_pytest.log
tail -3 gpurun_out/r02_call2_pytest.log
timeout 300 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest "tests/test_gpu_parity.py::test_full_physics_fused_on_structured_meshes[3-33-17-steps]" -m gpu -q -x > gpurun_out/r02_call2_memcheck.log 2>&1; echo "memcheck rc=$?" >> gpurun_out/r02_call2_memcheck.log
tail -4 gpurun_out/r02_call2_memcheck.log
python -m pytest "tests/test_fullsize_parity.py::test_full_mesh_rhs_against_oracle_and_arbiter[config2]" -m gpu -q -s 2>&1 | tail -15 > gpurun_out/r02_call2_fullsize_cfg2.log
tail -3 gpurun_out/r02_call2_fullsize_cfg2.log
python bench.py --mode full --no-cpu-baseline --steps 5 > gpurun_out/r02_bench_full_lserk4.json 2> gpurun_out/r02_bench_full_lserk4.err
python bench.py --mode full --stepper ssprk2 --no-cpu-baseline --steps 5 > gpurun_out/r02_bench_coastal_ssprk2.json 2> gpurun_out/r02_bench_coastal_ssprk2.err
python bench.py --mode full --variant 1 --no-cpu-baseline --no-e2e --no-parity --steps 5 > gpurun_out/r02_bench_full_lserk4_v1.json 2> /dev/null
cut -c1-200 gpurun_out/r02_bench_full_lserk4.json; cut -c1-200 gpurun_out/r02_bench_coastal_ssprk2.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r02_launches_full_lserk4.csv python bench.py --mode full --steps 1 --warmup 1 --reps 1 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_fused_persistent -s 20 -c 2 -o gpurun_out/r02_fused_full -f python bench.py --mode full --steps 1 --warmup 1 --reps 1 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/ncu_f.log 2>&1
ls -la gpurun_out | tail -12
