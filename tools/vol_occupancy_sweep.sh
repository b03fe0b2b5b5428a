for name in base vb4 vb5 vb6 base vb5; do
  if [ "$name" = base ]; then lib=""; else lib=$PWD/blitzdg_b200/csrc/libblitzdg_b200_$name.so; fi
  for a in "--mode full" "--variant 1"; do
  BDG_B200_LIB=$lib timeout 200 python bench.py $a --steps 3 --no-cpu-baseline --no-e2e 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$name', '$a', round(d['value']/1e9,2), 'G')"
  done
done
