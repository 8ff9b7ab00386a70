"""Op-level measurement of the on-device pre-processing (dclr_filter_clouds, dclr_subsample_clouds) at KITTI raw-scan
size against the reference's host path (numpy: SystematicErasing -> RangeSelection -> random choice, restated in
oracle/preproc.py).  One JSON line per op: kernel time (CUDA events, inputs larger than L2 or L2 flushed between
iterations), algorithmic bytes / time vs the measured HBM peak, and the numpy time per scan on one host core.
  python tools/preproc_bench.py [--scans 128] [--points 120000] > gpurun_out/preproc.jsonl"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deepclr_b200.data import device as dev_ops  # noqa: E402
from oracle import preproc  # noqa: E402


def hbm_peak():
    try:
        p = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
        for k in ('hbm_gbs', 'hbm_gbps'):
            if k in p:
                return float(p[k]), k
        for k, v in p.items():
            if 'hbm' in k.lower() and isinstance(v, (int, float)):
                return float(v), k
    except Exception:
        pass
    return 6550.0, 'fallback'


def timed(fn, iters, flush):
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    for a, b in ev:
        if flush is not None:
            flush.zero_()
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    return float(np.median([a.elapsed_time(b) for a, b in ev]))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--scans', type=int, default=128)
    ap.add_argument('--points', type=int, default=120000)
    ap.add_argument('--n-out', type=int, default=16384)
    ap.add_argument('--iters', type=int, default=20)
    a = ap.parse_args()
    dev = torch.device('cuda:0')
    peak, peak_src = hbm_peak()
    rng = np.random.Generator(np.random.PCG64(1234))
    sizes = rng.integers(int(a.points * 0.9), int(a.points * 1.1), a.scans).tolist()
    clouds = [np.concatenate([rng.uniform(-80, 80, (n, 2)), rng.uniform(-3, 3, (n, 2))], 1).astype(np.float32)
              for n in sizes]
    rc = dev_ops.RaggedClouds.from_list(clouds, dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    total = sum(sizes)

    for name, kw in (('filter nth=2 range[2,50]', dict(nth=2, start=0, min_range=2.0, max_range=50.0)),
                     ('filter range[2,50]', dict(nth=1, start=0, min_range=2.0, max_range=50.0))):
        out = dev_ops.filter_clouds(rc, **kw)
        kept = int(out.counts.sum())
        ms = timed(lambda: dev_ops.filter_clouds(rc, **kw), a.iters, flush)
        algo = 16 * total + 16 * kept
        t0 = time.perf_counter()
        for c in clouds[:8]:
            preproc.filter_cloud(c, kw['nth'], kw['start'], kw['min_range'], kw['max_range'])
        cpu_ms = (time.perf_counter() - t0) / 8 * 1e3
        print(json.dumps({'op': 'dclr_filter_clouds', 'case': name, 'scans': a.scans, 'points_in': total,
                          'points_kept': kept, 'ms': ms, 'algorithmic_bytes': algo,
                          'achieved_gbps': algo / ms / 1e6, 'peak_gbps': peak, 'peak_source': peak_src,
                          'frac': algo / ms / 1e6 / peak, 'l2': 'flushed (256 MB memset) before every launch',
                          'numpy_ms_per_scan_1core': cpu_ms, 'numpy_ms_all_scans_1core': cpu_ms * a.scans}))

    filt = dev_ops.filter_clouds(rc, nth=2, start=0, min_range=2.0, max_range=50.0)
    counts = filt.counts.cpu().numpy()
    n_out = min(a.n_out, int(counts.min()))
    ms = timed(lambda: dev_ops.subsample_clouds(filt, n_out, seed=1), a.iters, flush)
    # 16 B per chosen row read (a 32-byte sector is the least DRAM can move: 2x) + 16 B written
    algo = 2 * 16 * n_out * a.scans
    t0 = time.perf_counter()
    for b, c in enumerate(clouds[:8]):
        k = np.random.choice(np.arange(int(counts[b])), size=n_out, replace=False)
        c[:int(counts[b])][k, :]
    cpu_ms = (time.perf_counter() - t0) / 8 * 1e3
    print(json.dumps({'op': 'dclr_subsample_clouds', 'scans': a.scans, 'n_out': n_out, 'ms': ms,
                      'algorithmic_bytes': algo, 'achieved_gbps': algo / ms / 1e6, 'peak_gbps': peak,
                      'peak_source': peak_src, 'frac': algo / ms / 1e6 / peak,
                      'numpy_ms_per_scan_1core': cpu_ms, 'numpy_ms_all_scans_1core': cpu_ms * a.scans}))

    # whole front end from pinned host memory: H2D of the raw scans + filter + subsample (what replaces the host loop)
    stage = torch.cat([torch.from_numpy(c) for c in clouds]).pin_memory()
    off = rc.offsets

    def front():
        pts = stage.to(dev, non_blocking=True)
        r = dev_ops.filter_clouds(dev_ops.RaggedClouds(pts, off, sizes), nth=2, start=0, min_range=2.0, max_range=50.0)
        return dev_ops.subsample_clouds(r, n_out, seed=1)
    ms = timed(front, a.iters, None)
    print(json.dumps({'op': 'front end (pinned H2D + filter + subsample)', 'scans': a.scans, 'ms': ms,
                      'h2d_bytes': 16 * total, 'h2d_gbps': 16 * total / ms / 1e6, 'scans_per_s': a.scans / ms * 1e3}))


if __name__ == '__main__':
    main()
