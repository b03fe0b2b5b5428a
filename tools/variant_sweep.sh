#!/bin/bash
# Runs the short device-only bench for each libblitzdg_b200_<name>.so given (default build = "base").
#   tools/variant_sweep.sh base asmem prof
for name in "$@"; do
  if [ "$name" = base ]; then lib=""; else lib=$PWD/blitzdg_b200/csrc/libblitzdg_b200_$name.so; fi
  BDG_B200_LIB=$lib timeout 150 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-e2e 2> /tmp/err_$name.log | tail -1 | \
    python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$name', round(d['value']/1e9,2), 'G DSU/s frac', round(d['roofline']['frac'],4), d['clocks']['reasons'], 'eta', d['eta_max_after'])"
  grep PHASEPROF /tmp/err_$name.log
done
