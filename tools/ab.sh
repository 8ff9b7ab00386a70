#!/bin/bash
# A/B kernel experiments on ONE box: builds variants of the library with extra -D flags HERE (before gpurun), then
#   tools/ab.sh run             (on the GPU box) benches every deepclr_b200/lib_*.so and the default build, 2 rounds
#   tools/ab.sh build NAME -DFLAG...   (in the build container)
set -e
cd "$(dirname "$0")/.."
if [ "$1" = build ]; then
  name=$2; shift 2
  rm -rf /tmp/ab_$name && mkdir -p /tmp/ab_$name/pkg /tmp/ab_$name/include && cp -r deepclr_b200/csrc /tmp/ab_$name/pkg/csrc && cp include/*.h /tmp/ab_$name/include/
  rm -rf /tmp/ab_$name/pkg/csrc/build
  make -s -j8 -C /tmp/ab_$name/pkg/csrc NVFLAGS_EXTRA="$*" TARGET=../lib.so 2>&1 | grep -v "nvcc warning" || true
  cp /tmp/ab_$name/pkg/lib.so deepclr_b200/lib_$name.so
  echo "built deepclr_b200/lib_$name.so"
else
  for round in 1 2; do
    for lib in default deepclr_b200/lib_*.so; do
      [ "$lib" = default ] && unset DCLR_LIB || export DCLR_LIB=$PWD/$lib
      for shape in ${AB_SHAPES:-8}; do
      export DCLR_FPS_SHAPE=$shape
      python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-latency | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$lib', 'shape$shape', round(d['value']), {k: round(v, 4) for k, v in d['stage_ms'].items()})"
      done
    done
  done
fi
