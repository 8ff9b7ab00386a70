"""Timeline of the pose-head kernel's roles on CTA 0 (needs a -DDCLR_TRACE build:
   make -C deepclr_b200/csrc clean && make -C deepclr_b200/csrc NVFLAGS_EXTRA=-DDCLR_TRACE).  GPU only."""
import ctypes, os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from deepclr_b200 import _lib

model, _ = bench.build_model_and_cfg('kitti')
model = model.cuda()
model.numerics = 'fp16'
x = torch.from_numpy(bench.make_batches('kitti', 64, 16384, 1, 1)[0]).cuda()
with torch.no_grad():
    for _ in range(3):
        model(x)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * 1024)()
fn = _lib.lib().dclr_debug_read_trace
fn.argtypes = [ctypes.c_void_p, ctypes.c_int]
assert fn(buf, 1024) == 0
t = np.array(buf[:], dtype=np.int64)
for tile in range(4):
    b = tile * 64
    t0 = t[b + 16]
    if t0 == 0:
        continue
    rel = lambda i: int(t[b + i] - t0)
    print(f'tile {tile}: build {rel(16)}..{rel(17)}')
    for l in range(5):
        line = f'  L{l}: mma starts {rel(2 * l)}, issued by {rel(2 * l + 1)}'
        if l < 4:
            line += f' | epilogue {rel(18 + 2 * l)}..{rel(19 + 2 * l)}'
        print(line)
    print('  fin blocks done at', [rel(26 + nc) for nc in range(8)])

fn2 = _lib.lib().dclr_debug_read_flow_trace
fn2.argtypes = [ctypes.c_void_p, ctypes.c_int]
assert fn2(buf, 1024) == 0
f = np.array(buf[:128], dtype=np.int64).reshape(8, 16)
print('flow slot 0 of CTA 0, per tile: build | wait L1 | epi1 | wait L2 | epi2 | wait L3 | epi3 | total   (cycles)')
for i in range(8):
    x = f[i]
    if x[0] == 0:
        continue
    print(i, [int(x[j + 1] - x[j]) for j in range(7)], int(x[7] - x[0]), 'gap to next', int(f[i + 1][0] - x[7]) if i < 7 else '-')
    if x[8]:
        print('   build: record wait', int(x[11] - x[0]), '| coords+mask', int(x[8] - x[11]), '| wait for the gathered features',
              int(x[9] - x[8]), '| proxy fence', int(x[1] - x[9]))
