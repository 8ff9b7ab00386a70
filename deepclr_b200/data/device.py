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


# ---------------------------------------------------------------- pre-processing of raw scans (ragged batches)
class RaggedClouds:
    """Raw scans of different lengths back to back on the device: points [total, C] fp32, offsets [B+1] int32 (device),
    counts [B] int32 (device; rows in use per cloud, <= the slot length) and the host-known slot lengths."""

    def __init__(self, points: torch.Tensor, offsets: torch.Tensor, lengths, counts: Optional[torch.Tensor] = None):
        self.points, self.offsets, self.lengths, self.counts = points, offsets, list(lengths), counts

    @staticmethod
    def from_list(clouds, device, pinned: bool = True) -> 'RaggedClouds':
        """clouds: list of [n_i, C] float32 numpy arrays / CPU tensors (same C) -> one pinned staging buffer, one H2D."""
        ts = [torch.as_tensor(c, dtype=torch.float32) for c in clouds]
        if not ts:
            raise ValueError('no clouds')
        C = ts[0].shape[1]
        if any(t.dim() != 2 or t.shape[1] != C for t in ts):
            raise ValueError('all clouds must be [n, C] with the same C')
        lengths = [int(t.shape[0]) for t in ts]
        off = torch.zeros(len(ts) + 1, dtype=torch.int32)
        off[1:] = torch.cumsum(torch.tensor(lengths, dtype=torch.int64), 0).to(torch.int32)
        stage = torch.empty((int(off[-1]), C), dtype=torch.float32, pin_memory=pinned and torch.cuda.is_available())
        torch.cat(ts, 0, out=stage)
        return RaggedClouds(stage.to(device, non_blocking=True), off.to(device, non_blocking=True), lengths)

    @staticmethod
    def from_staged(stage: torch.Tensor, lengths, device) -> 'RaggedClouds':
        """stage: [total, C] float32 host tensor holding the clouds back to back (datasets.stage_scans)."""
        off = torch.zeros(len(lengths) + 1, dtype=torch.int32)
        off[1:] = torch.cumsum(torch.tensor(list(lengths), dtype=torch.int64), 0).to(torch.int32)
        return RaggedClouds(stage.to(device, non_blocking=True), off.to(device, non_blocking=True), lengths)

    def __len__(self) -> int:
        return len(self.lengths)


def filter_clouds(rc: RaggedClouds, nth: int = 1, start: int = 0, min_range: float = 0.0,
                  max_range: float = float('inf'), out_dim: Optional[int] = None) -> RaggedClouds:
    """SystematicErasing(nth, start) -> RangeSelection(min_range, max_range) -> TruncateDimension(out_dim)
    (deepclr/data/transforms/build.py:23-24,37-38) for every cloud of the ragged batch in one kernel; returns a ragged
    batch with the same offsets whose `counts` (device) hold the surviving rows per cloud."""
    _lib.require_cuda(rc.points, rc.offsets)
    pts = rc.points.contiguous()
    C = pts.shape[1]
    out_dim = C if out_dim is None else int(out_dim)
    if rc.counts is not None:
        raise RuntimeError('filter_clouds expects an unfiltered ragged batch (apply it once)')
    if nth == 1:
        start = 0
    out = torch.empty((pts.shape[0], out_dim), dtype=torch.float32, device=pts.device)
    counts = torch.empty(len(rc), dtype=torch.int32, device=pts.device)
    with torch.cuda.device(pts.device):
        _lib.check(_lib.lib().dclr_filter_clouds(_lib.ptr(pts), _lib.ptr(rc.offsets), len(rc), max(rc.lengths), C,
                                                 out_dim, int(nth), int(start), float(min_range), float(max_range),
                                                 _lib.ptr(out), _lib.ptr(counts), _lib.stream_ptr()),
                   'dclr_filter_clouds')
    return RaggedClouds(out, rc.offsets, rc.lengths, counts)


def subsample_clouds(rc: RaggedClouds, n_out: int, seed: int = 0, out_dim: Optional[int] = None,
                     ordered_if_equal: bool = True, return_perm: bool = False, return_short: bool = False):
    """n_out distinct random rows of every cloud -> dense [B, n_out, out_dim] (ModelInferenceHelper.stack,
    deepclr/models/base.py:130-135; RandomErasing max_points).  Optional extras: the chosen rows [B, n_out] int32 and
    the device counter of clouds that had fewer than n_out rows (their tail repeats rows)."""
    _lib.require_cuda(rc.points, rc.offsets)
    pts = rc.points.contiguous()
    C = pts.shape[1]
    out_dim = C if out_dim is None else int(out_dim)
    B = len(rc)
    x = torch.empty((B, int(n_out), out_dim), dtype=torch.float32, device=pts.device)
    perm = torch.empty((B, int(n_out)), dtype=torch.int32, device=pts.device) if return_perm else None
    short = torch.zeros(1, dtype=torch.int32, device=pts.device) if return_short else None
    with torch.cuda.device(pts.device):
        _lib.check(_lib.lib().dclr_subsample_clouds(_lib.ptr(pts), _lib.ptr(rc.offsets), _lib.ptr(rc.counts), B, C,
                                                    out_dim, int(n_out), int(seed) & (2 ** 64 - 1),
                                                    1 if ordered_if_equal else 0, _lib.ptr(x), _lib.ptr(perm),
                                                    _lib.ptr(short), _lib.stream_ptr()), 'dclr_subsample_clouds')
    extras = tuple(t for t in (perm, short) if t is not None)
    return (x,) + extras if extras else x


def prepare_pairs(templates, sources, n_points: int, device, input_dim: Optional[int] = None, nth: int = 1,
                  start: int = 0, min_range: float = 0.0, max_range: float = float('inf'), seed: int = 0) -> torch.Tensor:
    """Raw host scans of B pairs -> the forward's input x [2B, n_points, input_dim] on the device, templates first
    (deepclr/data/build.py:82): one upload, one filter kernel, one subsample kernel."""
    if len(templates) != len(sources):
        raise ValueError('need as many templates as sources')
    rc = RaggedClouds.from_list(list(templates) + list(sources), device)
    rc = filter_clouds(rc, nth, start, min_range, max_range, out_dim=input_dim)
    return subsample_clouds(rc, n_points, seed=seed)
