"""FPS kernel variants side by side on one box: ms per launch of dclr_fps (C ABI) on KITTI-shaped 16384-point clouds,
for warps-per-cloud x samples-per-round (DCLR_FPS_SHAPE x DCLR_FPS_MP, read per call) and several batch sizes."""
import ctypes
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deepclr_b200 import _lib, synth  # noqa: E402

pool = np.stack([synth.kitti_pair(1234 + i, 16384)[i % 2] for i in range(16)])
variants = [(8, 0), (8, 1), (8, 2), (8, 3), (8, 4), (8, 6), (16, 2), (16, 4)]
if len(sys.argv) > 1:         # e.g. 8x3 (warps x samples per round), c3 = cluster of two CTAs per cloud, 3 samples per round
    variants = [(-2, int(a[1:])) if a[0] == 'c' else tuple(int(v) for v in a.split('x')) for a in sys.argv[1:]]
ref = None
for G in (128, 32, 16, 2):
    x = torch.from_numpy(np.ascontiguousarray(pool[np.arange(G) % 16])).cuda()
    x[:, :, :3] += torch.rand(G, 1, 3, device='cuda') * 0.05
    idx = torch.empty((G, 1024), dtype=torch.int32, device='cuda')
    nx = torch.empty((G, 1024, 3), dtype=torch.float32, device='cuda')
    base = None
    for shape, mp in variants:
        os.environ['DCLR_FPS_SHAPE'], os.environ['DCLR_FPS_MP'] = str(shape if shape > 0 else 8), str(mp)
        os.environ['DCLR_FPS_CLUSTER'] = '1' if shape < 0 else '0' 
        call = lambda: _lib.check(_lib.lib().dclr_fps(_lib.ptr(x), 4, G, 16384, 1024, 1, _lib.ptr(idx), _lib.ptr(nx),
                                                      _lib.stream_ptr()), 'fps')
        for _ in range(3):
            call()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            call()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        got = idx.cpu().numpy()
        if base is None:
            base = got
        print(json.dumps({'clouds': G, 'warps': shape, 'mp': mp, 'ms': round(ms, 4), 'same_as_first': bool((got == base).all())}),
              flush=True)
