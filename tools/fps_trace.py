"""Per-round phase timing of the bucketed FPS kernel, CTA 0, rounds 500..531 (needs a -DDCLR_TRACE build)."""
import ctypes, os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deepclr_b200 import _lib, synth
from deepclr_b200.pointnet2 import pointnet2_utils as pu
xyz = torch.from_numpy(np.stack([synth.kitti_pair(1234 + i, 16384)[0][:, :3] for i in range(4)])).cuda().contiguous()
for _ in range(3):
    pu.furthest_point_sample(xyz, 1024)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * 4096)()
fn = _lib.lib().dclr_debug_read_fps_trace
fn.argtypes = [ctypes.c_void_p, ctypes.c_int]
assert fn(buf, 4096) == 0
t = np.array(buf[:4096], dtype=np.int64).reshape(32, 8, 16)
print('round: per-warp [updates] | warp0: test, upd, reduce, barrier+final | round length (warp0 start to next start)')
for r in range(31):
    w = t[r]
    ups = [int(x) for x in w[:, 6]]
    crit = int(np.argmax(w[:, 2] - w[:, 1]))
    print(r, ups, '| w0:', int(w[0, 1] - w[0, 0]), int(w[0, 2] - w[0, 1]), int(w[0, 3] - w[0, 2]), int(w[0, 4] - w[0, 3]),
          f'| slowest warp {crit}: upd {int(w[crit, 2] - w[crit, 1])} arrives {int(w[crit, 3] - w[0, 0])}',
          '| round', int(t[r + 1, 0, 0] - w[0, 0]))
print('single-update rounds (warp with exactly 1 update): ballot->math, switch, redux1, redux2, ->end of updates')
for r in range(31):
    for w in range(8):
        if t[r, w, 6] == 1:
            x = t[r, w]
            print(r, w, int(x[8] - x[1]), int(x[9] - x[8]), int(x[10] - x[9]), int(x[11] - x[10]), int(x[2] - x[11]))
print('mean round', float(np.mean(t[1:31, 0, 0] - t[0:30, 0, 0])))
