"""Per-round phase timing of the bucketed FPS kernels, CTA 0 (needs a -DDCLR_TRACE build: tools/ab.sh build trace -DDCLR_TRACE,
then DCLR_LIB=deepclr_b200/lib_trace.so python tools/fps_trace.py).  DCLR_FPS_MP=1: the single-pick kernel, rounds
500..531; otherwise the multi-pick kernel, rounds 100..131 (slots: 0 start, 1 after skip test, 2 after updates, 3 entry
published, 4 picks known; 6 = buckets updated, 7 = picks applied this round)."""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deepclr_b200 import _lib, synth  # noqa: E402
from deepclr_b200.pointnet2 import pointnet2_utils as pu  # noqa: E402

xyz = torch.from_numpy(np.stack([synth.kitti_pair(1234 + i, 16384)[0][:, :3] for i in range(4)])).cuda().contiguous()
for _ in range(3):
    pu.furthest_point_sample(xyz, 1024)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * 4096)()
mp = int(os.environ.get('DCLR_FPS_MP', '4'))
fn = _lib.lib().dclr_debug_read_fps_trace if mp == 0 else _lib.lib().dclr_debug_read_fps_trace_tm
fn.argtypes = [ctypes.c_void_p, ctypes.c_int]
assert fn(buf, 4096) == 0
t = np.array(buf[:4096], dtype=np.int64).reshape(32, 8, 16)
if mp == 0:
    print('round: per-warp [updates] | warp0: test, upd, reduce, barrier+final | round length (warp0 start to next start)')
    for r in range(31):
        w = t[r]
        ups = [int(x) for x in w[:, 6]]
        crit = int(np.argmax(w[:, 2] - w[:, 1]))
        print(r, ups, '| w0:', int(w[0, 1] - w[0, 0]), int(w[0, 2] - w[0, 1]), int(w[0, 3] - w[0, 2]), int(w[0, 4] - w[0, 3]),
              f'| slowest warp {crit}: upd {int(w[crit, 2] - w[crit, 1])} arrives {int(w[crit, 3] - w[0, 0])}',
              '| round', int(t[r + 1, 0, 0] - w[0, 0]))
    print('mean round', float(np.mean(t[1:31, 0, 0] - t[0:30, 0, 0])))
else:
    print('round picks [buckets updated per warp] | skip test, slowest update (warp), slowest entry, last arrival -> picks known | round')
    rounds = []
    for r in range(31):
        w = t[r]
        ups = [int(x) for x in w[:, 6]]
        start = int(w[:, 0].min())
        crit = int(np.argmax(w[:, 2] - w[:, 1]))
        last_arr = int(w[:, 3].max())
        rl = int(t[r + 1, :, 0].min() - start)
        rounds.append(rl)
        print(r, int(w[0, 7]), ups, '| test', int(np.median(w[:, 1] - w[:, 0])), f'upd {int(w[crit, 2] - w[crit, 1])} (w{crit})',
              'entry', int((w[:, 3] - w[:, 2]).max()), 'arrive', last_arr - start, 'select', int(w[:, 4].max() - last_arr), '| round', rl)
    print('mean round', float(np.mean(rounds)), 'mean picks', float(t[:31, 0, 7].mean()),
          'mean updates per round', float(t[:31, :, 6].sum(1).mean()))
