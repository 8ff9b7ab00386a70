#!/bin/bash
# One GPU visit: parity tests, bench (both arms), ModelNet-shaped config, op sweep, ncu launch list, ncu --set full
# of the hot kernels.  Results land in gpurun_out/; tools/ncu_summary.py <tag> turns them into profiles/.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
cat gpurun_out/bench_ref.json
python bench.py --kind modelnet --points 1024 --pairs 1024 --steps 20 --warmup 5 --cpu-pairs 64 > gpurun_out/bench_modelnet.json 2> gpurun_out/bench_modelnet.err; echo "modelnet rc=$?"
cat gpurun_out/bench_modelnet.json
python bench.py --numerics bf16 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_bf16.json 2>/dev/null; echo "bf16 rc=$?"
python bench.py --numerics fp32 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_fp32.json 2>/dev/null; echo "fp32 rc=$?"
python tools/op_sweep.py > gpurun_out/op_sweep.jsonl 2> gpurun_out/op_sweep.err; echo "sweep rc=$?"
: > gpurun_out/b1_reference_cuda.jsonl; for cfg in "1 0" "1 1" "8 0" "8 1"; do set -- $cfg; python tools/bench_ref_cuda.py --pairs $1 --tf32 $2 >> gpurun_out/b1_reference_cuda.jsonl 2>> gpurun_out/b1.err; done; echo "b1 rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-latency > gpurun_out/ncu_launch.log 2>&1; echo "ncu list rc=$?"
for k in ${NCU_KERNELS:-fps_bucket_kernel}; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 6 -c 1 -f -o gpurun_out/full_$k \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-latency > gpurun_out/ncu_full_$k.log 2>&1; echo "ncu full $k rc=$?"
done
# pre-processing kernels: op bench + ncu --set full of both
timeout 300 python tools/preproc_bench.py > gpurun_out/preproc.jsonl 2> gpurun_out/preproc.err; echo "preproc bench rc=$?"
cat gpurun_out/preproc.jsonl
for k in filter_clouds_kernel subsample_clouds_kernel; do
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:$k -s 4 -c 1 -f -o gpurun_out/full_$k \
    python tools/preproc_bench.py --iters 2 > gpurun_out/ncu_full_$k.log 2>&1; echo "ncu full $k rc=$?"
done
