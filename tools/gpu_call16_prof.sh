#!/bin/bash
mkdir -p gpurun_out
BDG_B200_LIB=$PWD/blitzdg_b200/csrc/libblitzdg_b200_prof.so timeout 120 python bench.py --steps 3 --warmup 3 --reps 1 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/r02_phaseprof.json 2> gpurun_out/r02_phaseprof.err
grep PHASEPROF gpurun_out/r02_phaseprof.err
python -c "
import json; d=json.loads(open('gpurun_out/r02_phaseprof.json').read().strip().split('\n')[-1]); print(round(d['value']/1e9,2), d['clocks']['sm_mhz'])"
