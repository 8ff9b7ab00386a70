"""CPU fp32 restatement of the reference's registration forward path.  TEST INFRASTRUCTURE ONLY.

Functional (no nn.Module) restatement driven by a reference-keyed state dict and the
`params` block of a reference model_config.yaml.  Point ops come from oracle/ops.py (C restatement of the
reference kernels), dense layers are plain torch CPU fp32 matmuls.  Cited lines are in /root/reference:

  set abstraction   deepclr/models/deepclr.py:88-94, extern/Pointnet2.PyTorch/pointnet2/pointnet2_modules.py:19-55,
                    pointnet2_utils.py:241-264 (QueryAndGroup), pytorch_utils.py:5-32 (SharedMLP)
  motion embedding  deepclr/models/deepclr.py:157-173 (KnnGrouping), :201-231 (MotionEmbeddingBase)
  output            deepclr/models/deepclr.py:284-294, helper.py:11-38,41-65
  model             deepclr/models/deepclr.py:488-521

oracle/gen_golden.py checks this file against the reference's own Python modules imported from /root/reference
(possible only in the build container) and writes tests/golden/*.npz; that is what pins it.
"""
from typing import Dict, Optional

import numpy as np
import torch

from . import ops

LABEL_DIMS = {'POSE3D_EULER': 6, 'POSE3D_QUAT': 7, 'POSE3D_DUAL_QUAT': 8}


def _pointwise(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """1x1 conv + bias + ReLU over channel dim 1 of x [B,Cin,...]; w [Cout,Cin,1(,1)]."""
    w2 = w.reshape(w.shape[0], w.shape[1])
    y = torch.einsum('oc,bc...->bo...', w2, x) + b.reshape(1, -1, *([1] * (x.dim() - 2)))
    return torch.relu(y)


def fps_fn_default(xyz: np.ndarray, npoint: int) -> np.ndarray:
    return ops.fps_lowidx(xyz, npoint)


def sa_msg(sd: Dict[str, torch.Tensor], prefix: str, xyz: torch.Tensor, feats: Optional[torch.Tensor],
           npoint: int, radii, nsamples, fps_fn=fps_fn_default, debug: Optional[dict] = None):
    """One PointnetSAModuleMSG. xyz [B,N,3], feats [B,C,N] or None -> new_xyz [B,npoint,3], [B,sum(Cout),npoint]."""
    xyz_np = xyz.numpy()
    fidx = fps_fn(xyz_np, npoint)                                   # pointnet2_modules.py:34
    new_xyz_np = ops.gather(xyz_np.transpose(0, 2, 1), fidx).transpose(0, 2, 1).copy()   # :32-35
    new_xyz = torch.from_numpy(new_xyz_np)
    outs = []
    for s, (r, ns) in enumerate(zip(radii, nsamples)):
        idx, cnt = ops.ball_query(r, ns, xyz_np, new_xyz_np, return_cnt=True)   # pointnet2_utils.py:249
        g_xyz = torch.from_numpy(ops.group(xyz_np.transpose(0, 2, 1), idx))     # :250-251
        g_xyz = g_xyz - new_xyz.transpose(1, 2).unsqueeze(-1)                   # :252
        if feats is not None:
            g_f = torch.from_numpy(ops.group(feats.numpy(), idx))               # :255
            h = torch.cat([g_xyz, g_f], dim=1)                                  # :257 (xyz channels first)
        else:
            h = g_xyz
        layer = 0
        while f'{prefix}.mlps.{s}.layer{layer}.conv.weight' in sd:
            h = _pointwise(h, sd[f'{prefix}.mlps.{s}.layer{layer}.conv.weight'],
                           sd[f'{prefix}.mlps.{s}.layer{layer}.conv.bias'])
            layer += 1
        outs.append(h.max(dim=3).values)                                        # pointnet2_modules.py:41-52
        if debug is not None:
            debug[f'ball_idx{s}'] = idx
            debug[f'ball_cnt{s}'] = cnt
    if debug is not None:
        debug['fps_idx'] = fidx
    return new_xyz, torch.cat(outs, dim=1)


def cloud_features(sd, params: dict, x: torch.Tensor, fps_fn=fps_fn_default, debug: Optional[dict] = None):
    """x [2B,N,C] -> [2B, 3+F, npoint]  (deepclr.py:510-521, :88-94)."""
    p = params['cloud_features']['params']
    xyz = x[:, :, :3].contiguous()
    feats = x[:, :, 3:].transpose(1, 2).contiguous() if x.shape[2] > 3 else None
    for level in range(len(p['npoint'])):
        xyz, feats = sa_msg(sd, f'_cloud_layers.0._sa{level}', xyz, feats, p['npoint'][level], p['radii'][level],
                            p['nsamples'][level], fps_fn=fps_fn, debug=debug if level == 0 else None)
    return torch.cat([xyz.transpose(1, 2), feats], dim=1)


def motion_embedding(sd, params: dict, feat: torch.Tensor, debug: Optional[dict] = None) -> torch.Tensor:
    """feat [2B,D,n] (templates first) -> [B, 3+Cout, n]  (deepclr.py:201-231, 243-246)."""
    p = params['merge']['params']
    k, radius = int(p['k']), float(p['radius'])
    append = bool(p.get('append_features', True))
    B = feat.shape[0] // 2
    n, D = feat.shape[2], feat.shape[1]
    pts0 = feat[:B].transpose(1, 2).reshape(B * n, D)
    pts1 = feat[B:].transpose(1, 2).reshape(B * n, D)
    if k == 0:                                                    # GlobalGrouping, deepclr.py:108-139
        row = torch.arange(B * n).view(-1, 1).repeat(1, n)
        col = torch.arange(B * n).view(B, n).repeat(1, n).view(B * n, n)
    else:
        gi = ops.knn(pts1[:, :3].numpy(), pts0[:, :3].numpy(), k, B)   # deepclr.py:164-167
        gi = torch.from_numpy(gi).view(2, B * n, k)
        row, col = gi[0], gi[1]
    g0, g1 = pts0[row], pts1[col]                                 # :170-171
    pos_diff = g1[:, :, :3] - g0[:, :, :3]
    if append:
        merged = torch.cat([pos_diff, g0[:, :, 3:], g1[:, :, 3:]], dim=2)
    else:
        merged = torch.cat([pos_diff, g1[:, :, 3:] - g0[:, :, 3:]], dim=2)
    h = merged.transpose(1, 2)
    layer = 0
    base = '_merge_layers.0._embedding._conv._sequential'
    while f'{base}.{layer}._sequential.0.weight' in sd:
        h = _pointwise(h, sd[f'{base}.{layer}._sequential.0.weight'], sd[f'{base}.{layer}._sequential.0.bias'])
        layer += 1
    if radius > 0.0:                                              # :220-223
        mask = torch.norm(pos_diff, dim=2) >= radius
        h = torch.where(mask.unsqueeze(1), torch.zeros_like(h), h)
    f = h.max(dim=2).values                                       # :225
    out = torch.cat([pts0[:, :3], f], dim=1)
    if debug is not None and k > 0:
        debug['knn_idx'] = gi.numpy()
    return out.view(B, n, -1).transpose(1, 2).contiguous()


def output_simple(sd, label_type: str, x: torch.Tensor) -> torch.Tensor:
    """x [B,D,n] -> y [B,label_dim]  (deepclr.py:284-294, 275-282)."""
    h = x
    layer = 0
    while f'_merge_layers.1.conv._sequential.{layer}._sequential.0.weight' in sd:
        h = _pointwise(h, sd[f'_merge_layers.1.conv._sequential.{layer}._sequential.0.weight'],
                       sd[f'_merge_layers.1.conv._sequential.{layer}._sequential.0.bias'])
        layer += 1
    h = h.max(dim=2).values
    layer = 0
    while f'_merge_layers.1.linear._sequential.{layer}._sequential.0.weight' in sd:
        w = sd[f'_merge_layers.1.linear._sequential.{layer}._sequential.0.weight']
        b = sd[f'_merge_layers.1.linear._sequential.{layer}._sequential.0.bias']
        h = torch.relu(h @ w.t() + b)
        layer += 1
    y = h @ sd['_merge_layers.1.output.weight'].t() + sd['_merge_layers.1.output.bias']
    y = y.clone()
    if label_type == 'POSE3D_QUAT':
        y[:, 3] = torch.sigmoid(y[:, 3])
        y[:, 4:] = torch.tanh(y[:, 4:])
    elif label_type == 'POSE3D_DUAL_QUAT':
        y[:, 0] = torch.sigmoid(y[:, 0])
        y[:, 1:4] = torch.tanh(y[:, 1:4])
    return y


def forward(sd, cfg: dict, x: torch.Tensor, is_feat: bool = False, fps_fn=fps_fn_default,
            debug: Optional[dict] = None) -> torch.Tensor:
    """Whole model: x [2B,N,C] (or SA features if is_feat) -> y [B,label_dim]  (deepclr.py:488-508)."""
    sd = {k: v.detach().to(torch.float32).cpu() for k, v in sd.items()}
    x = x.detach().to(torch.float32).cpu()
    with torch.no_grad():
        feat = x if is_feat else cloud_features(sd, cfg['params'], x, fps_fn=fps_fn, debug=debug)
        emb = motion_embedding(sd, cfg['params'], feat, debug=debug)
        y = output_simple(sd, str(cfg['label_type']).upper(), emb)
    if debug is not None:
        debug['cloud_feat'] = feat.numpy()
        debug['motion_feat'] = emb.numpy()
    return y


# ---- pose comparison (labels.py:46-51,93-99; evaluation/metrics.py:8-49), transforms3d restated in numpy ----

def _quat2mat(q):
    w, x, y, z = q
    nq = w * w + x * x + y * y + z * z
    if nq < np.finfo(np.float64).eps:
        return np.eye(3)
    s = 2.0 / nq
    X, Y, Z = x * s, y * s, z * s
    wX, wY, wZ = w * X, w * Y, w * Z
    xX, xY, xZ = x * X, x * Y, x * Z
    yY, yZ, zZ = y * Y, y * Z, z * Z
    return np.array([[1.0 - (yY + zZ), xY - wZ, xZ + wY],
                     [xY + wZ, 1.0 - (xX + zZ), yZ - wX],
                     [xZ - wY, yZ + wX, 1.0 - (xX + yY)]])


def _qmult(a, b):
    w1, x1, y1, z1 = a
    w2, x2, y2, z2 = b
    return np.array([w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                     w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2, w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2])


def dual_quat_to_matrix(label) -> np.ndarray:
    label = np.asarray(label, dtype=np.float64)
    real, dual = label[:4], label[4:]
    nrm = np.sqrt(np.dot(real, real)) + 1e-8
    real, dual = real / nrm, dual / nrm
    m = np.eye(4)
    m[:3, :3] = _quat2mat(real)
    t = 2.0 * _qmult(dual, real * np.array([1.0, -1.0, -1.0, -1.0]))
    m[:3, 3] = t[1:]
    return m


def pose_errors(y_a, y_b):
    """(rotation error [deg], translation error [m]) between two dual-quat labels, KITTI definition."""
    d = np.linalg.inv(dual_quat_to_matrix(y_a)) @ dual_quat_to_matrix(y_b)
    c = max(min(0.5 * (d[0, 0] + d[1, 1] + d[2, 2] - 1.0), 1.0), -1.0)
    return float(np.rad2deg(np.arccos(c))), float(np.linalg.norm(d[:3, 3]))
