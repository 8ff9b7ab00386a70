"""Autograd-function facade over the C ABI, same names/layouts as the reference's pointnet2_utils.py
(furthest_point_sample :10-36, gather_operation :39-73, grouping_operation :156-197, ball_query :200-228,
QueryAndGroup :231-264).  The reference allocates outputs in Python and hands raw pointers to pybind wrappers;
the same ownership rule applies here with ctypes."""
from typing import Optional, Tuple

import torch
import torch.nn as nn
from torch.autograd import Function

from .. import _lib


def _tie_rule_id() -> int:
    import deepclr_b200
    return deepclr_b200.FPS_TIE_RULES[deepclr_b200.get_fps_tie_rule()]


def _call(name, *args):
    _lib.check(getattr(_lib.lib(), name)(*args, _lib.stream_ptr()), name)


class FurthestPointSampling(Function):
    @staticmethod
    def forward(ctx, xyz: torch.Tensor, npoint: int) -> torch.Tensor:
        """xyz (B, N, 3) -> (B, npoint) int32 indices; first index 0, ties by deepclr_b200.get_fps_tie_rule()."""
        assert xyz.is_contiguous()
        _lib.require_cuda(xyz)
        B, N, _ = xyz.size()
        output = torch.empty((B, npoint), dtype=torch.int32, device=xyz.device)
        with torch.cuda.device(xyz.device):
            _call('dclr_fps', _lib.ptr(xyz), 3, B, N, npoint, _tie_rule_id(), _lib.ptr(output), None)
        ctx.mark_non_differentiable(output)
        return output

    @staticmethod
    def backward(ctx, a=None):
        return None, None


furthest_point_sample = FurthestPointSampling.apply


class GatherOperation(Function):
    @staticmethod
    def forward(ctx, features: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
        """features (B, C, N), idx (B, npoint) int32 -> (B, C, npoint)."""
        assert features.is_contiguous()
        assert idx.is_contiguous()
        _lib.require_cuda(features, idx)
        B, npoint = idx.size()
        _, C, N = features.size()
        output = torch.empty((B, C, npoint), dtype=torch.float32, device=features.device)
        with torch.cuda.device(features.device):
            _call('dclr_gather', _lib.ptr(features), _lib.ptr(idx), B, C, N, npoint, _lib.ptr(output))
        ctx.save_for_backward(idx)
        ctx.n = N
        return output

    @staticmethod
    def backward(ctx, grad_out):
        # scatter-add, as the reference's gather_points_grad kernel (pointnet2_utils.py:60-73, sampling_gpu.cu:46-63)
        idx, = ctx.saved_tensors
        B, C, M = grad_out.shape
        grad = torch.zeros((B, C, ctx.n), dtype=torch.float32, device=grad_out.device)
        g = grad_out.contiguous().float()
        with torch.cuda.device(grad_out.device):
            _call('dclr_gather_grad', _lib.ptr(g), _lib.ptr(idx), B, C, ctx.n, M, _lib.ptr(grad))
        return grad, None


gather_operation = GatherOperation.apply


class GroupingOperation(Function):
    @staticmethod
    def forward(ctx, features: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
        """features (B, C, N), idx (B, npoint, nsample) int32 -> (B, C, npoint, nsample)."""
        assert features.is_contiguous()
        assert idx.is_contiguous()
        _lib.require_cuda(features, idx)
        B, nfeatures, nsample = idx.size()
        _, C, N = features.size()
        output = torch.empty((B, C, nfeatures, nsample), dtype=torch.float32, device=features.device)
        with torch.cuda.device(features.device):
            _call('dclr_group', _lib.ptr(features), _lib.ptr(idx), B, C, N, nfeatures, nsample, _lib.ptr(output))
        ctx.save_for_backward(idx)
        ctx.n = N
        return output

    @staticmethod
    def backward(ctx, grad_out: torch.Tensor) -> Tuple[torch.Tensor, None]:
        # scatter-add, as the reference's group_points_grad kernel (pointnet2_utils.py:181-197, group_points_gpu.cu:8-25)
        idx, = ctx.saved_tensors
        B, C, M, ns = grad_out.shape
        grad = torch.zeros((B, C, ctx.n), dtype=torch.float32, device=grad_out.device)
        g = grad_out.contiguous().float()
        with torch.cuda.device(grad_out.device):
            _call('dclr_group_grad', _lib.ptr(g), _lib.ptr(idx), B, C, ctx.n, M, ns, _lib.ptr(grad))
        return grad, None


grouping_operation = GroupingOperation.apply


class BallQuery(Function):
    @staticmethod
    def forward(ctx, radius: float, nsample: int, xyz: torch.Tensor, new_xyz: torch.Tensor) -> torch.Tensor:
        """xyz (B, N, 3), new_xyz (B, npoint, 3) -> (B, npoint, nsample) int32, padded like the reference."""
        assert new_xyz.is_contiguous()
        assert xyz.is_contiguous()
        _lib.require_cuda(xyz, new_xyz)
        B, N, _ = xyz.size()
        npoint = new_xyz.size(1)
        idx = torch.empty((B, npoint, nsample), dtype=torch.int32, device=xyz.device)
        with torch.cuda.device(xyz.device):
            _call('dclr_ball_query', _lib.ptr(new_xyz), _lib.ptr(xyz), 3, B, N, npoint, float(radius), nsample,
                  _lib.ptr(idx), None, 1)
        ctx.mark_non_differentiable(idx)
        return idx

    @staticmethod
    def backward(ctx, a=None):
        return None, None, None, None


ball_query = BallQuery.apply


class QueryAndGroup(nn.Module):
    def __init__(self, radius: float, nsample: int, use_xyz: bool = True):
        super().__init__()
        self.radius, self.nsample, self.use_xyz = radius, nsample, use_xyz

    def forward(self, xyz: torch.Tensor, new_xyz: torch.Tensor, features: Optional[torch.Tensor] = None):
        """-> (B, 3 + C, npoint, nsample), relative xyz channels first."""
        idx = ball_query(self.radius, self.nsample, xyz, new_xyz)
        grouped = grouping_operation(xyz.transpose(1, 2).contiguous(), idx)
        grouped = grouped - new_xyz.transpose(1, 2).unsqueeze(-1)
        if features is None:
            assert self.use_xyz, "Cannot have not features and not use xyz as a feature!"
            return grouped
        grouped_features = grouping_operation(features, idx)
        return torch.cat([grouped, grouped_features], dim=1) if self.use_xyz else grouped_features
