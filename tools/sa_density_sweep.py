"""Grouped shared-MLP + max-pool against neighbourhood density, both kernels (sa_mlp_kernel: fp32 SIMT; sa_tc_kernel:
tcgen05 with fp16 hi/lo split operands; DCLR_SA_ROUTE selects; genuine rows only): synthetic
clouds whose balls hold ~4 .. ~1000 points (points uniform in a cube, radii chosen for the target counts, nsample =
512 / 1024 as in the KITTI config), through the stage-level C entry dclr_sa_msg_fused; rows really computed (sum of the
ball counts), rows/s, actual fp32 FLOP/s and the fraction of the FFMA peak (148 SMs x 128 lanes x 2 x SM clock).
One JSON line per density.  python tools/sa_density_sweep.py > gpurun_out/sa_density.jsonl"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

dev = torch.device('cuda:0')
model, cfg = bench.build_model_and_cfg('kitti')
model = model.to(dev)
model.numerics = 'fp16'
G, N, M = 32, 16384, 1024
rng = np.random.default_rng(1)
with torch.no_grad():
    model(torch.from_numpy(bench.make_batches('kitti', 1, N, 1, 1)[0]).to(dev))
native = model._native
r0, r1 = 0.5, 1.0                                  # the model's radii (KITTI config); density is set by the cube size
flop_row = 2.0 * (4 * 16 + 16 * 16 + 16 * 32)
clock = 1.965e9
peak = 148 * 128 * 2 * clock / 1e12
for target in (1, 2, 4, 8, 16, 32, 128, 512, 1000):
    # expected count in the SMALL ball = N * (4/3 pi r0^3) / L^3  ->  cube edge L
    L = (N * 4.0 / 3.0 * np.pi * r0 ** 3 / target) ** (1.0 / 3.0)
    pts = rng.uniform(0, L, (G, N, 4)).astype(np.float32)
    x = torch.from_numpy(pts).to(dev)
    centres = x[:, :M, :3].contiguous()
    for route in ('simt', 'tc'):
      os.environ['DCLR_SA_ROUTE'] = route
      with torch.no_grad():
        native.sa_msg(x, centres)
        cnt = [native.debug_tensor(1 + s, (G, M)).clamp(min=1) for s in range(2)]
        rows = int(cnt[0].sum().item() + cnt[1].sum().item())
        native.set_timing(True)
        for _ in range(10):
            native.sa_msg(x, centres)
        torch.cuda.synchronize()
        st, calls = native.collect_timing()
        native.set_timing(False)
      ms = st['sa_mlp_maxpool'] / calls
      print(json.dumps({'kernel': 'sa_mlp_kernel' if route == 'simt' else 'sa_tc_kernel', 'target_rows_small_ball': target, 'mean_rows_per_centre': [float(c.float().mean().item()) for c in cnt],
                      'clouds': G, 'centres_per_cloud': M, 'rows_computed': rows, 'ms': ms, 'Grows_per_s': rows / ms / 1e6,
                      'actual_TFLOPs_fp32': rows * flop_row / ms / 1e9, 'ffma_peak_TFLOPs': peak,
                      'frac_of_ffma_peak': rows * flop_row / ms / 1e9 / peak,
                      'ball_query_ms': st['ball_query'] / calls}), flush=True)
