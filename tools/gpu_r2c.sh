#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "fps" > gpurun_out/r2c_pytest_fps.log 2>&1; tail -3 gpurun_out/r2c_pytest_fps.log
python tools/fps_ab.py > gpurun_out/r2c_fps_ab.jsonl 2> gpurun_out/r2c_fps_ab.err; head -9 gpurun_out/r2c_fps_ab.jsonl; tail -9 gpurun_out/r2c_fps_ab.jsonl
for v in "8 4" "8 2"; do set -- $v
  DCLR_LIB=$PWD/deepclr_b200/lib_trace.so DCLR_FPS_SHAPE=$1 DCLR_FPS_MP=$2 python tools/fps_trace.py > gpurun_out/r2c_fps_trace_$1x$2.txt 2>&1
  tail -6 gpurun_out/r2c_fps_trace_$1x$2.txt
done
nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/micro/lat3 tools/micro/lat3.cu && tools/micro/lat3 > gpurun_out/r2b_lat3.txt 2>&1; cat gpurun_out/r2b_lat3.txt
ncu --set full --import-source on --clock-control none -k regex:fps_tmem_kernel -c 1 -f -o gpurun_out/r2c_fps_tmem_8x4 python tools/fps_ab.py 8x4 > gpurun_out/r2c_ncu.log 2>&1
