"""Sequential (odometry) mode, SURVEY.md 8f rank 1: per-frame latency of `ModelInferenceHelper(is_sequential=True)`
(1 set abstraction + 1 merge per frame, features of frame t-1 cached on the device; deepclr/models/base.py:97-112,
scripts/inference.py:83-119) -- the time the paper's "Time [ms]" column measures.  Frames are synthetic HDL-64-like
scans: 16384 points (BASELINE.json's size) and ~60 k points (real KITTI after SystematicErasing(2),
scripts/converter/kitti_odometry.py:14), the latter also from RAW ~120 k-point scans with the device-side filter in
front.  CUDA events around predict() as in scripts/inference.py:95-110.  One JSON line per case.
    python tools/seq_bench.py > gpurun_out/seq.jsonl"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import build_model_and_cfg  # noqa: E402
from deepclr_b200 import synth  # noqa: E402
from deepclr_b200.data import device as dev_ops  # noqa: E402
from deepclr_b200.models import ModelInferenceHelper  # noqa: E402


def scan(rng, n):
    xyz, inten = synth._lidar_scan(rng, n)
    return np.concatenate([xyz, inten[:, None]], 1).astype(np.float32)


def run(helper, frames, prep=None):
    ts = []
    helper.reset_state()
    for i, f in enumerate(frames):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        x = prep(f) if prep else f
        y = helper.predict(x)
        b.record()
        b.synchronize()
        if i >= 5:
            assert y is not None and torch.isfinite(y).all()
            ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2], ts[int(len(ts) * 0.9)]


def main():
    dev = torch.device('cuda:0')
    model, _ = build_model_and_cfg('kitti')
    model = model.to(dev)
    model.numerics = os.environ.get('DCLR_NUMERICS', 'fp16')
    helper = ModelInferenceHelper(model, is_sequential=True)
    rng = np.random.Generator(np.random.PCG64(99))
    for n in (16384, 60000):
        frames = [torch.from_numpy(scan(rng, n)).to(dev) for _ in range(35)]
        p50, p90 = run(helper, frames)
        print(json.dumps({'mode': 'sequential', 'points_per_frame': n, 'input': 'device-resident frames',
                          'p50_ms': p50, 'p90_ms': p90, 'frames': 30}), flush=True)
        graphed = ModelInferenceHelper(model, is_sequential=True, use_cuda_graph=True)
        p50, p90 = run(graphed, frames)
        print(json.dumps({'mode': 'sequential, per-frame step replayed from one CUDA graph', 'points_per_frame': n,
                          'input': 'device-resident frames', 'p50_ms': p50, 'p90_ms': p90, 'frames': 30}), flush=True)
    raw = [torch.from_numpy(scan(rng, int(rng.integers(115000, 125000)))).pin_memory() for _ in range(35)]

    def prep(host):
        rc = dev_ops.RaggedClouds.from_staged(host, [host.shape[0]], dev)
        rc = dev_ops.filter_clouds(rc, nth=2, start=0, min_range=2.0, max_range=80.0)
        n = int(rc.counts[0])                      # the one host read-back of the frame: its surviving length
        return rc.points[:n]
    p50, p90 = run(helper, raw, prep)
    print(json.dumps({'mode': 'sequential', 'points_per_frame': '~120000 raw -> every 2nd point, range [2, 80] m',
                      'input': 'pinned raw scans (H2D + dclr_filter_clouds inside the timed region)',
                      'p50_ms': p50, 'p90_ms': p90, 'frames': 30}), flush=True)


if __name__ == '__main__':
    main()
