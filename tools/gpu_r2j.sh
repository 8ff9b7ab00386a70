#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q -k "fps_multipick" > gpurun_out/r2j_pytest_fps.log 2>&1; tail -4 gpurun_out/r2j_pytest_fps.log
timeout 300 python tools/fps_ab.py 8x3 c2 c3 c4 c6 > gpurun_out/r2j_fps_ab.jsonl 2> gpurun_out/r2j_fps_ab.err; cat gpurun_out/r2j_fps_ab.jsonl; tail -3 gpurun_out/r2j_fps_ab.err
