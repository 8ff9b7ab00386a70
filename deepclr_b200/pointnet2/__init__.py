"""Drop-in for the `pointnet2` package the reference installs from extern/Pointnet2.PyTorch
(as patched by extern/pointnet2.patch:7-9,18,32,49), restricted to what DeepCLR's forward path uses."""
from .pointnet2_modules import PointnetSAModuleMSG, PointnetSAModule
from .pointnet2_utils import (furthest_point_sample, gather_operation, ball_query, grouping_operation,
                              QueryAndGroup)
from .pytorch_utils import SharedMLP

__all__ = ['PointnetSAModuleMSG', 'PointnetSAModule', 'furthest_point_sample', 'gather_operation', 'ball_query',
           'grouping_operation', 'QueryAndGroup', 'SharedMLP']
