"""Pose deviation of the 16-bit tensor-core modes next to the deviation the REFERENCE shows against itself when torch's
default TF32 convolutions are on (torch.backends.cudnn.allow_tf32 = True is the default since torch 1.7; the reference pins
torch 1.8.1): both measured against the reference composition in strict fp32 (reference CUDA kernels from oracle/_ref +
torch fp32 layers), on full-size KITTI-shaped pairs with the seeded random-init weights.  One JSON line."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from oracle import model_oracle, ref_cuda  # noqa: E402

pairs = int(sys.argv[1]) if len(sys.argv) > 1 else 16
model, cfg = bench.build_model_and_cfg('kitti')
model = model.cuda()
x = torch.from_numpy(bench.make_batches('kitti', pairs, 16384, 1, 4242)[0]).cuda()
sd = model.state_dict()


def ref(tf32_conv, tf32_mm):
    torch.backends.cudnn.allow_tf32 = tf32_conv
    torch.backends.cuda.matmul.allow_tf32 = tf32_mm
    return ref_cuda.forward(sd, cfg, x).cpu().numpy()


def errs(a, b):
    e = [model_oracle.pose_errors(a[i], b[i]) for i in range(a.shape[0])]
    return {'rot_deg_max': max(r for r, _ in e), 'rot_deg_mean': float(np.mean([r for r, _ in e])),
            'trans_m_max': max(t for _, t in e), 'trans_m_mean': float(np.mean([t for _, t in e]))}


y32 = ref(False, False)
out = {'pairs': pairs, 'reference_tf32_conv_default_vs_reference_fp32': errs(ref(True, False), y32),
       'reference_tf32_conv_and_matmul_vs_reference_fp32': errs(ref(True, True), y32)}
for mode in ('fp32', 'fp16', 'bf16'):
    model.numerics = mode
    with torch.no_grad():
        y = model(x)[0].cpu().numpy()
    out[f'this_repo_{mode}_vs_reference_fp32'] = errs(y, y32)
print(json.dumps(out))
