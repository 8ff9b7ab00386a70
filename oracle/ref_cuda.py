"""The REFERENCE's own CUDA kernels (oracle/_ref/libpointnet2_ref.so, built by oracle/build_ref.sh from the sources in
/root/reference, unmodified) behind ctypes, plus the reference forward composed from them on the GPU.
TEST INFRASTRUCTURE / BASELINE ONLY (second oracle for the point ops on the GPU box; baseline B1 of BASELINE.md).

Wrappers follow extern/Pointnet2.PyTorch/pointnet2/pointnet2_utils.py of the reference: the caller allocates the outputs
exactly as the autograd Functions do (:25-28 temp filled with 1e10, :55-57, :172-174, :218-220 idx zeroed).
kNN: torch_cluster is not vendored, so `knn_standin` (cdist^2 + stable top-k) stands in for it -- flagged wherever it
is used.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, '_ref', 'libpointnet2_ref.so')
_lib = None

_SYMS = {
    'fps': '_Z39furthest_point_sampling_kernel_launcheriiiPKfPfPiP11CUstream_st',
    'gather': '_Z34gather_points_kernel_launcher_fastiiiiPKfPKiPfP11CUstream_st',
    'ball': '_Z31ball_query_kernel_launcher_fastiiifiPKfS0_PiP11CUstream_st',
    'group': '_Z33group_points_kernel_launcher_fastiiiiiPKfPKiPfP11CUstream_st',
    'gather_grad': '_Z39gather_points_grad_kernel_launcher_fastiiiiPKfPKiPfP11CUstream_st',
    'group_grad': '_Z38group_points_grad_kernel_launcher_fastiiiiiPKfPKiPfP11CUstream_st',
}


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(LIB_PATH)
        vp, i, f = ctypes.c_void_p, ctypes.c_int, ctypes.c_float
        getattr(_lib, _SYMS['fps']).argtypes = [i, i, i, vp, vp, vp, vp]
        getattr(_lib, _SYMS['gather']).argtypes = [i, i, i, i, vp, vp, vp, vp]
        getattr(_lib, _SYMS['ball']).argtypes = [i, i, i, f, i, vp, vp, vp, vp]
        getattr(_lib, _SYMS['group']).argtypes = [i, i, i, i, i, vp, vp, vp, vp]
        getattr(_lib, _SYMS['gather_grad']).argtypes = [i, i, i, i, vp, vp, vp, vp]
        getattr(_lib, _SYMS['group_grad']).argtypes = [i, i, i, i, i, vp, vp, vp, vp]
        for s in _SYMS.values():
            getattr(_lib, s).restype = None
    return _lib


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


def _st():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def furthest_point_sample(xyz: torch.Tensor, npoint: int) -> torch.Tensor:
    assert xyz.is_cuda and xyz.is_contiguous()
    B, N, _ = xyz.shape
    idx = torch.empty((B, npoint), dtype=torch.int32, device=xyz.device)
    temp = torch.full((B, N), 1e10, dtype=torch.float32, device=xyz.device)
    getattr(lib(), _SYMS['fps'])(B, N, npoint, _p(xyz), _p(temp), _p(idx), _st())
    return idx


def gather_operation(features: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    assert features.is_contiguous() and idx.is_contiguous()
    B, C, N = features.shape
    M = idx.shape[1]
    out = torch.empty((B, C, M), dtype=torch.float32, device=features.device)
    getattr(lib(), _SYMS['gather'])(B, C, N, M, _p(features), _p(idx), _p(out), _st())
    return out


def ball_query(radius: float, nsample: int, xyz: torch.Tensor, new_xyz: torch.Tensor) -> torch.Tensor:
    assert xyz.is_contiguous() and new_xyz.is_contiguous()
    B, N, _ = xyz.shape
    M = new_xyz.shape[1]
    idx = torch.zeros((B, M, nsample), dtype=torch.int32, device=xyz.device)
    getattr(lib(), _SYMS['ball'])(B, N, M, ctypes.c_float(radius), nsample, _p(new_xyz), _p(xyz), _p(idx), _st())
    return idx


def grouping_operation(features: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    assert features.is_contiguous() and idx.is_contiguous()
    B, C, N = features.shape
    _, M, ns = idx.shape
    out = torch.empty((B, C, M, ns), dtype=torch.float32, device=features.device)
    getattr(lib(), _SYMS['group'])(B, C, N, M, ns, _p(features), _p(idx), _p(out), _st())
    return out


def gather_operation_grad(grad_out: torch.Tensor, idx: torch.Tensor, N: int) -> torch.Tensor:
    """pointnet2_utils.py:60-73: grad_features zero-filled by the caller, then gather_points_grad_wrapper."""
    B, C, M = grad_out.shape
    grad = torch.zeros((B, C, N), dtype=torch.float32, device=grad_out.device)
    getattr(lib(), _SYMS['gather_grad'])(B, C, N, M, _p(grad_out.contiguous()), _p(idx), _p(grad), _st())
    return grad


def grouping_operation_grad(grad_out: torch.Tensor, idx: torch.Tensor, N: int) -> torch.Tensor:
    """pointnet2_utils.py:181-197: group_points_grad_wrapper(B, C, N, npoint, nsample, grad_out, idx, grad_features)."""
    B, C, M, ns = grad_out.shape
    grad = torch.zeros((B, C, N), dtype=torch.float32, device=grad_out.device)
    getattr(lib(), _SYMS['group_grad'])(B, C, N, M, ns, _p(grad_out.contiguous()), _p(idx), _p(grad), _st())
    return grad


def knn_standin(x: torch.Tensor, y: torch.Tensor, k: int, B: int) -> torch.Tensor:
    """STAND-IN for torch_cluster.knn (absent): x [B*Nx,3] candidates, y [B*Ny,3] queries -> [2, B*Ny*k] int64."""
    Nx, Ny = x.shape[0] // B, y.shape[0] // B
    d = torch.cdist(y.view(B, Ny, 3), x.view(B, Nx, 3)) ** 2
    col = torch.topk(d, k, dim=2, largest=False, sorted=True).indices
    col = col + (torch.arange(B, device=x.device) * Nx).view(B, 1, 1)
    row = torch.arange(B * Ny, device=x.device).view(B, Ny, 1).expand(-1, -1, k)
    return torch.stack([row.reshape(-1), col.reshape(-1)], 0)


def _pointwise(x, w, b):
    """1x1 conv + bias + ReLU (pytorch_utils.py:21-32 / helper.py:11-38) through torch's conv -> cuDNN/cuBLAS fp32."""
    if x.dim() == 4:
        return torch.relu_(torch.nn.functional.conv2d(x, w.reshape(w.shape[0], w.shape[1], 1, 1), b))
    return torch.relu_(torch.nn.functional.conv1d(x, w.reshape(w.shape[0], w.shape[1], 1), b))


def set_abstraction(sd, cfg: dict, x: torch.Tensor) -> torch.Tensor:
    """SetAbstraction.forward (deepclr.py:88-94 -> pointnet2_modules.py:19-55): x [G,N,C] -> [G,3+F,npoint]."""
    sa = cfg['params']['cloud_features']['params']
    with torch.no_grad():
        xyz = x[:, :, :3].contiguous()                                             # deepclr.py:30-37
        feats = x[:, :, 3:].transpose(1, 2).contiguous() if x.shape[2] > 3 else None
        xyz_flipped = xyz.transpose(1, 2).contiguous()                             # pointnet2_modules.py:30
        fidx = furthest_point_sample(xyz, sa['npoint'][0])
        new_xyz = gather_operation(xyz_flipped, fidx).transpose(1, 2).contiguous()
        outs = []
        for s, (r, ns) in enumerate(zip(sa['radii'][0], sa['nsamples'][0])):
            idx = ball_query(r, ns, xyz, new_xyz)                                  # pointnet2_utils.py:249
            g = grouping_operation(xyz_flipped, idx) - new_xyz.transpose(1, 2).unsqueeze(-1)
            if feats is not None:
                g = torch.cat([g, grouping_operation(feats, idx)], dim=1)
            layer = 0
            while f'_cloud_layers.0._sa0.mlps.{s}.layer{layer}.conv.weight' in sd:
                g = _pointwise(g, sd[f'_cloud_layers.0._sa0.mlps.{s}.layer{layer}.conv.weight'],
                               sd[f'_cloud_layers.0._sa0.mlps.{s}.layer{layer}.conv.bias'])
                layer += 1
            outs.append(torch.nn.functional.max_pool2d(g, kernel_size=[1, g.size(3)]).squeeze(-1))
        return torch.cat([new_xyz.transpose(1, 2), torch.cat(outs, dim=1)], dim=1)     # [G,67,M]


def motion_embedding(sd, cfg: dict, feat: torch.Tensor) -> torch.Tensor:
    """MotionEmbedding.forward (deepclr.py:157-246): feat [2B,3+F,n] templates first -> [B,3+C,n]; kNN = stand-in."""
    m = cfg['params']['merge']['params']
    with torch.no_grad():
        k, radius = int(m['k']), float(m['radius'])
        B, D, n = feat.shape[0] // 2, feat.shape[1], feat.shape[2]
        pts0 = feat[:B].transpose(1, 2).reshape(B * n, D)
        pts1 = feat[B:].transpose(1, 2).reshape(B * n, D)
        gi = knn_standin(pts1[:, :3].contiguous(), pts0[:, :3].contiguous(), k, B).view(2, B * n, k)
        g0, g1 = pts0[gi[0]], pts1[gi[1]]
        pos_diff = g1[:, :, :3] - g0[:, :, :3]
        h = torch.cat([pos_diff, g0[:, :, 3:], g1[:, :, 3:]], dim=2).transpose(1, 2)
        base = '_merge_layers.0._embedding._conv._sequential'
        layer = 0
        while f'{base}.{layer}._sequential.0.weight' in sd:
            h = _pointwise(h, sd[f'{base}.{layer}._sequential.0.weight'], sd[f'{base}.{layer}._sequential.0.bias'])
            layer += 1
        if radius > 0.0:
            h = h.masked_fill((torch.norm(pos_diff, dim=2) >= radius).unsqueeze(1), 0.0)
        f = h.max(dim=2).values
        return torch.cat([pts0[:, :3], f], dim=1).view(B, n, -1).transpose(1, 2).contiguous()


def output_head(sd, emb: torch.Tensor) -> torch.Tensor:
    """OutputSimple.forward (deepclr.py:284-294), dual-quaternion activation (:279-281): emb [B,3+C,n] -> [B,8]."""
    with torch.no_grad():
        h = emb
        layer = 0
        while f'_merge_layers.1.conv._sequential.{layer}._sequential.0.weight' in sd:
            h = _pointwise(h, sd[f'_merge_layers.1.conv._sequential.{layer}._sequential.0.weight'],
                           sd[f'_merge_layers.1.conv._sequential.{layer}._sequential.0.bias'])
            layer += 1
        h = h.max(dim=2).values
        layer = 0
        while f'_merge_layers.1.linear._sequential.{layer}._sequential.0.weight' in sd:
            h = torch.relu(torch.nn.functional.linear(
                h, sd[f'_merge_layers.1.linear._sequential.{layer}._sequential.0.weight'],
                sd[f'_merge_layers.1.linear._sequential.{layer}._sequential.0.bias']))
            layer += 1
        y = torch.nn.functional.linear(h, sd['_merge_layers.1.output.weight'], sd['_merge_layers.1.output.bias'])
        return torch.cat([torch.sigmoid(y[:, 0:1]), torch.tanh(y[:, 1:4]), y[:, 4:]], dim=1)


def forward(sd, cfg: dict, x: torch.Tensor) -> torch.Tensor:
    """Reference forward (deepclr.py:488-521) on the GPU: reference CUDA ops + torch fp32 dense layers.
    x [2B,N,C] cuda -> y [B,8]."""
    return output_head(sd, motion_embedding(sd, cfg, set_abstraction(sd, cfg, x)))
