"""PointnetSAModuleMSG with the constructor, forward signature and parameter names of the reference
(pointnet2_modules.py:19-55, :58-92).  This is the op-by-op composition (FPS -> gather -> per scale ball query,
group, SharedMLP, max-pool) on the C-ABI ops; DeepCLR's whole-model forward uses the fused C entry instead
(deepclr_b200/models/deepclr.py)."""
from typing import List, Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import pointnet2_utils
from . import pytorch_utils as pt_utils


class PointnetSAModuleMSG(nn.Module):
    def __init__(self, *, npoint: int, radii: List[float], nsamples: List[int], mlps: List[List[int]],
                 bn: bool = True, use_xyz: bool = True, pool_method: str = 'max_pool', instance_norm: bool = False):
        super().__init__()
        assert len(radii) == len(nsamples) == len(mlps)
        if instance_norm:
            raise NotImplementedError('instance_norm is not used by DeepCLR')
        if npoint is None:
            raise NotImplementedError('GroupAll (npoint=None) is not used by DeepCLR')
        self.npoint = npoint
        self.radii, self.nsamples = list(radii), list(nsamples)
        self.pool_method = pool_method
        self.groupers = nn.ModuleList()
        self.mlps = nn.ModuleList()
        for radius, nsample, spec in zip(radii, nsamples, mlps):
            self.groupers.append(pointnet2_utils.QueryAndGroup(radius, nsample, use_xyz=use_xyz))
            spec = list(spec)
            if use_xyz:
                spec[0] += 3
            self.mlps.append(pt_utils.SharedMLP(spec, bn=bn))

    def forward(self, xyz: torch.Tensor, features: Optional[torch.Tensor] = None,
                new_xyz: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        """xyz (B,N,3), features (B,C,N) -> new_xyz (B,npoint,3), new_features (B, sum(mlps[k][-1]), npoint)."""
        if new_xyz is None:
            idx = pointnet2_utils.furthest_point_sample(xyz, self.npoint)
            new_xyz = pointnet2_utils.gather_operation(xyz.transpose(1, 2).contiguous(), idx)\
                .transpose(1, 2).contiguous()
        pooled = []
        for grouper, mlp in zip(self.groupers, self.mlps):
            h = mlp(grouper(xyz, new_xyz, features))
            if self.pool_method == 'max_pool':
                h = F.max_pool2d(h, kernel_size=[1, h.size(3)])
            elif self.pool_method == 'avg_pool':
                h = F.avg_pool2d(h, kernel_size=[1, h.size(3)])
            else:
                raise NotImplementedError
            pooled.append(h.squeeze(-1))
        return new_xyz, torch.cat(pooled, dim=1)


class PointnetSAModule(PointnetSAModuleMSG):
    def __init__(self, *, mlp: List[int], npoint: int = None, radius: float = None, nsample: int = None,
                 bn: bool = True, use_xyz: bool = True, pool_method: str = 'max_pool', instance_norm: bool = False):
        super().__init__(mlps=[mlp], npoint=npoint, radii=[radius], nsamples=[nsample], bn=bn, use_xyz=use_xyz,
                         pool_method=pool_method, instance_norm=instance_norm)
