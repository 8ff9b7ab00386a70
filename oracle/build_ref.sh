#!/usr/bin/env bash
# Builds the REFERENCE's own CUDA kernels for sm_100a from the sources where they lie under /root/reference
# (extern/Pointnet2.PyTorch/pointnet2/src/{sampling,ball_query,group_points}_gpu.cu, unmodified, never copied) into
# oracle/_ref/libpointnet2_ref.so.  TEST INFRASTRUCTURE / BASELINE ONLY: used by tests (second oracle for the point
# ops) and by tools/bench_ref_cuda.py (baseline B1 "reference CUDA ops on B200").  Only the raw-pointer launchers
# (*_kernel_launcher*, declared in src/*_gpu.h) are used; the THC-era pybind wrappers are not built.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
SRC=/root/reference/extern/Pointnet2.PyTorch/pointnet2/src
OUT="$HERE/_ref"
[ -d "$SRC" ] || { echo "reference tree not present; keeping prebuilt $OUT" >&2; exit 0; }
mkdir -p "$OUT"
if [ -f "$OUT/libpointnet2_ref.so" ] && [ "$OUT/libpointnet2_ref.so" -nt "$SRC/sampling_gpu.cu" ]; then exit 0; fi
PY=${PYTHON:-python}
INC=$($PY -c "import torch.utils.cpp_extension as c, sysconfig; print(' '.join('-I'+p for p in c.include_paths()+[sysconfig.get_paths()['include']]))")
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
pids=()
for n in sampling ball_query group_points; do
  $NVCC -O2 -std=c++17 -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC $INC -c "$SRC/${n}_gpu.cu" -o "$OUT/${n}_gpu.o" &
  pids+=($!)
done
for p in "${pids[@]}"; do wait "$p"; done
$NVCC -shared -o "$OUT/libpointnet2_ref.so" "$OUT"/sampling_gpu.o "$OUT"/ball_query_gpu.o "$OUT"/group_points_gpu.o -lcudart
rm -f "$OUT"/*.o
echo "built $OUT/libpointnet2_ref.so"
