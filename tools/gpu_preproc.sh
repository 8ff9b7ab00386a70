#!/bin/bash
# GPU visit for the pre-processing kernels: their parity tests first (own timeout: the filter kernel spins on a chain
# word), then the whole GPU suite, the op bench, a short bench line and ncu --set full of the two kernels.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_preproc.py -x -q > gpurun_out/pytest_preproc.log 2>&1; echo "preproc pytest rc=$?"
tail -15 gpurun_out/pytest_preproc.log
timeout 300 python tools/preproc_bench.py > gpurun_out/preproc.jsonl 2> gpurun_out/preproc.err; echo "preproc bench rc=$?"
cat gpurun_out/preproc.jsonl; tail -3 gpurun_out/preproc.err
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
cat gpurun_out/bench_quick.json
for k in filter_clouds_kernel subsample_clouds_kernel; do
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:$k -s 4 -c 1 -f -o gpurun_out/full_$k \
    python tools/preproc_bench.py --iters 2 > gpurun_out/ncu_full_$k.log 2>&1; echo "ncu full $k rc=$?"
done
