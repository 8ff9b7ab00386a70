"""Batched, on-device counterparts of the per-pair host post-processing of the reference's inference loop
(scripts/inference.py:113-119): label -> 4x4 matrix (LabelType.to_matrix, deepclr/data/labels.py:80-101) and the KITTI
rotation / translation errors (deepclr/evaluation/metrics.py:8-49).  float64 on the GPU through the C ABI
(dclr_pose_to_matrix / dclr_pose_errors); poses never leave the device to be compared."""
from typing import Optional, Tuple

import torch

from .. import _lib
from .labels import LabelType


def to_matrix(labels: torch.Tensor, label_type: LabelType, scale: Optional[float] = None) -> torch.Tensor:
    """labels [B, label_type.dim] fp32 CUDA -> [B, 4, 4] float64 CUDA."""
    _lib.require_cuda(labels)
    if labels.dim() != 2 or labels.shape[1] != label_type.dim:
        raise ValueError(f'labels must be [B,{label_type.dim}], got {tuple(labels.shape)}')
    y = labels.contiguous().float()
    out = torch.empty((y.shape[0], 4, 4), dtype=torch.float64, device=y.device)
    with torch.cuda.device(y.device):
        _lib.check(_lib.lib().dclr_pose_to_matrix(_lib.ptr(y), y.shape[0], label_type.native_id,
                                                  float(scale) if scale else 0.0, _lib.ptr(out), _lib.stream_ptr()),
                   'dclr_pose_to_matrix')
    return out


def kitti_errors(m1: torch.Tensor, m2: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """m1, m2 [B,4,4] float64 CUDA -> (rotation error [rad], translation error) per pair, KITTI definition."""
    _lib.require_cuda(m1, m2)
    if m1.shape != m2.shape or m1.shape[1:] != (4, 4):
        raise ValueError('expected two [B,4,4] tensors')
    a, b = m1.contiguous().double(), m2.contiguous().double()
    rot = torch.empty(a.shape[0], dtype=torch.float64, device=a.device)
    trans = torch.empty_like(rot)
    with torch.cuda.device(a.device):
        _lib.check(_lib.lib().dclr_pose_errors(_lib.ptr(a), _lib.ptr(b), a.shape[0], _lib.ptr(rot), _lib.ptr(trans),
                                               _lib.stream_ptr()), 'dclr_pose_errors')
    return rot, trans
