"""Translation / rotation label losses used by the in-model loss modules (reference deepclr/utils/metrics.py:40-141).
Pure torch; training is outside the accelerated path."""
from typing import Optional

import torch

from ..data.labels import LabelType


def _reduce(x: torch.Tensor, reduction: Optional[str]) -> torch.Tensor:
    if reduction is None or reduction == 'none':
        return x
    if reduction == 'mean':
        return torch.mean(x)
    if reduction == 'sum':
        return torch.sum(x)
    raise RuntimeError(f"Unsupported reduction '{reduction}'")


def _normalize(x: torch.Tensor, label_type: LabelType, eps: float) -> torch.Tensor:
    if label_type == LabelType.POSE3D_QUAT:
        return torch.cat((x[:, :3], x[:, 3:] / (torch.norm(x[:, 3:], p=2, dim=1, keepdim=True) + eps)), dim=1)
    if label_type == LabelType.POSE3D_DUAL_QUAT:
        return x / (torch.norm(x[:, :4], p=2, dim=1, keepdim=True) + eps)
    raise RuntimeError("Unsupported label type for normalization")


def trans_loss(source, target, label_type: LabelType, p: int = 2, reduction: Optional[str] = 'mean',
               eps: float = 1e-8) -> torch.Tensor:
    if label_type == LabelType.POSE3D_DUAL_QUAT:
        a, b = _normalize(source, label_type, eps)[:, 4:], _normalize(target, label_type, eps)[:, 4:]
    else:
        a, b = source[:, :3], target[:, :3]
    return _reduce(torch.norm(a - b, dim=1, p=p, keepdim=True), reduction)


def rot_loss(source, target, label_type: LabelType, p: int = 2, reduction: Optional[str] = 'mean',
             eps: float = 1e-8) -> torch.Tensor:
    if label_type == LabelType.POSE3D_EULER:
        a, b = source[:, 3:], target[:, 3:]
    elif label_type == LabelType.POSE3D_QUAT:
        a, b = _normalize(source, label_type, eps)[:, 3:], _normalize(target, label_type, eps)[:, 3:]
    else:
        a, b = _normalize(source, label_type, eps)[:, :4], _normalize(target, label_type, eps)[:, :4]
    return _reduce(torch.norm(a - b, dim=1, p=p, keepdim=True), reduction)
