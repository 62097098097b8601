#!/bin/bash
# tools/variant_bench.sh NAME...  : C2 kernel time of each variant library, both layouts
for V in "$@"; do
  if [ "$V" = base ]; then unset RBX_LIBRARY; else export RBX_LIBRARY=$PWD/randblas_b200/variants/librbx_$V.so; fi
  for L in R C; do
    timeout 120 python bench.py --steps 5 --warmup 3 --layout $L --no-extras --no-cpu --e2e-steps 0 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$V', d['config'].get('layout'), round(d['ms_per_step'],3), round(d['roofline']['frac'],4))"
  done
done
