#!/bin/bash
# source-level ncu capture of the FPS sampling kernel (variant from $1, e.g. 8x4) + micro latencies
mkdir -p gpurun_out
V=${1:-8x4}
ncu --set full --import-source on --clock-control none -k regex:fps_tmem_kernel -c 1 -f -o gpurun_out/r2b_fps_tmem_$V python tools/fps_ab.py $V > gpurun_out/r2b_ncu_$V.log 2>&1
tail -3 gpurun_out/r2b_ncu_$V.log
if [ -f tools/micro/lat3.cu ]; then nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/micro/lat3 tools/micro/lat3.cu && tools/micro/lat3 > gpurun_out/r2b_lat3.txt 2>&1; cat gpurun_out/r2b_lat3.txt; fi
