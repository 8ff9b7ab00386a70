"""Baseline B1: the reference's CUDA ops recompiled for sm_100a (oracle/_ref) + torch fp32 dense layers,
i.e. the reference forward as it would run on a B200, timed with the reference's protocol (CUDA events around the
forward, scripts/timing.py:32-47) on the same synthetic KITTI-shaped pairs as bench.py.  kNN is a torch stand-in
(torch_cluster is not vendored).  Prints one JSON line."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
from oracle import ref_cuda

ap = argparse.ArgumentParser()
ap.add_argument('--pairs', type=int, default=8)
ap.add_argument('--points', type=int, default=16384)
ap.add_argument('--kind', default='kitti')
ap.add_argument('--steps', type=int, default=20)
ap.add_argument('--warmup', type=int, default=5)
ap.add_argument('--tf32', type=int, default=0)
args = ap.parse_args()
torch.backends.cuda.matmul.allow_tf32 = bool(args.tf32)
torch.backends.cudnn.allow_tf32 = bool(args.tf32)
model, cfg = bench.build_model_and_cfg(args.kind)
sd = {k: v.cuda() for k, v in model.state_dict().items()}
xs = [torch.from_numpy(b).cuda() for b in bench.make_batches(args.kind, args.pairs, args.points, 4, 1234)]
for i in range(args.warmup):
    ref_cuda.forward(sd, cfg, xs[i % 4])
torch.cuda.synchronize()
times = []
for i in range(args.steps):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    ref_cuda.forward(sd, cfg, xs[i % 4])
    b.record()
    b.synchronize()
    times.append(a.elapsed_time(b))
times.sort()
mean = sum(times) / len(times)
print(json.dumps({'baseline': 'B1 reference CUDA ops (sm_100a rebuild) + torch fp32 dense, kNN stand-in',
                  'kind': args.kind, 'pairs_per_step': args.pairs, 'points': args.points, 'tf32': bool(args.tf32),
                  'ms_per_step_mean': mean, 'ms_p50': times[len(times) // 2],
                  'pairs_per_s': args.pairs / (mean * 1e-3), 'peak_mem_gb': torch.cuda.max_memory_allocated() / 2**30}))
