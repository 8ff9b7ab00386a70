#!/bin/bash
# A/B of the filter kernel's CTAs per SM (register budget): default build + deepclr_b200/lib_*.so
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_preproc.py -x -q 2>&1 | tail -3
for lib in default deepclr_b200/lib_*.so; do
  [ "$lib" = default ] && unset DCLR_LIB || export DCLR_LIB=$PWD/$lib
  for round in 1 2; do
  timeout 300 python tools/preproc_bench.py 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); print('$lib', d['op'][:22], d.get('case',''), round(d['ms'],4), round(d.get('achieved_gbps',0)))"
  done
done | tee gpurun_out/preproc_ab.txt
