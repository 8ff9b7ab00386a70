"""Drop-in for `torch_cluster.knn` as used by the reference (deepclr/models/deepclr.py:164-166)."""
from typing import Optional

import torch

from . import _lib


def knn(x: torch.Tensor, y: torch.Tensor, k: int, batch_x: Optional[torch.Tensor] = None,
        batch_y: Optional[torch.Tensor] = None, cosine: bool = False, num_workers: int = 1) -> torch.Tensor:
    """For every row of y the k nearest rows of x within the same batch element -> LongTensor [2, Ny*k]
    (row 0 = y index, row 1 = x index, ascending distance, ties -> lower x index).

    Batches must be sorted and equal-sized, which is what the reference builds (deepclr.py:150-155)."""
    if cosine:
        raise NotImplementedError('cosine distance is not used by DeepCLR and not built')
    _lib.require_cuda(x, y)
    if x.dim() != 2 or y.dim() != 2 or x.shape[1] != 3 or y.shape[1] != 3:
        raise ValueError('knn expects [N,3] inputs')
    B = 1
    if batch_x is not None:
        B = int(batch_x[-1].item()) + 1 if batch_x.numel() else 1
    if x.shape[0] % B or y.shape[0] % B:
        raise ValueError('knn: batches must be equal-sized')
    if batch_x is not None and batch_y is not None and x.shape[0] and y.shape[0]:
        nx, ny = x.shape[0] // B, y.shape[0] // B
        ok = bool((batch_x.view(B, nx) == torch.arange(B, device=x.device).view(B, 1)).all()) and \
            bool((batch_y.view(B, ny) == torch.arange(B, device=y.device).view(B, 1)).all())
        if not ok:
            raise ValueError('knn: batch vectors must be sorted with equal-sized elements')
    x = x.contiguous().float()
    y = y.contiguous().float()
    out = torch.empty((2, y.shape[0] * k), dtype=torch.long, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(_lib.lib().dclr_knn(_lib.ptr(x), _lib.ptr(y), B, x.shape[0] // B, y.shape[0] // B, k,
                                       _lib.ptr(out[0]), _lib.ptr(out[1]), _lib.stream_ptr()), 'dclr_knn')
    return out
