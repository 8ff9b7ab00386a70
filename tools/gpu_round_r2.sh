#!/bin/bash
# One GPU visit for the round-2 profile set: parity tests, smoke, bench (both arms), ModelNet-shaped configs, other
# numerics, op sweep, sequential mode, ncu launch list, ncu --set full of the hot kernels.  Results land in gpurun_out/;
# `python tools/ncu_summary.py r2` turns them into profiles/.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/smoke.log
python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
python bench.py --kind modelnet --points 1024 --pairs 1024 --steps 20 --warmup 5 --cpu-pairs 64 > gpurun_out/bench_modelnet.json 2> gpurun_out/bench_modelnet.err; echo "modelnet rc=$?"
python bench.py --kind modelnet --points 1024 --pairs 1024 --partial --noise 0.04 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_modelnet_partial.json 2>/dev/null; echo "modelnet partial rc=$?"
python bench.py --numerics bf16 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_bf16.json 2>/dev/null; echo "bf16 rc=$?"
python bench.py --numerics fp32 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_fp32.json 2>/dev/null; echo "fp32 rc=$?"
python bench.py --streams 1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_streams1.json 2>/dev/null; echo "streams1 rc=$?"
python tools/op_sweep.py > gpurun_out/op_sweep.jsonl 2> gpurun_out/op_sweep.err; echo "sweep rc=$?"
python tools/seq_bench.py > gpurun_out/seq.jsonl 2> gpurun_out/seq.err; echo "seq rc=$?"
python tools/fps_ab.py 8x0 8x1 8x2 8x3 8x4 8x6 16x4 > gpurun_out/fps_ab.jsonl 2>/dev/null; echo "fps ab rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv \
  python bench.py --streams 1 --steps 2 --warmup 3 --no-cpu-baseline --no-latency --no-raw-scans > gpurun_out/ncu_launch.log 2>&1; echo "ncu list rc=$?"
for k in ${NCU_KERNELS:-fps_tmem_kernel flow_tc_kernel head_tc_kernel knn_fast_kernel bq_grid_query_kernel sa_mlp_kernel fps_sort_kernel bq_grid_build_kernel linear_skinny_kernel}; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 6 -c 1 -f -o gpurun_out/full_$k \
    python bench.py --streams 1 --steps 2 --warmup 3 --no-cpu-baseline --no-latency --no-raw-scans > gpurun_out/ncu_full_$k.log 2>&1; echo "ncu full $k rc=$?"
done
