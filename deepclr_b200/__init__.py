"""deepclr_b200: B200-native (sm_100a) implementation of DeepCLR's registration forward path.

Host-side mirror of the reference interface for this path:
  deepclr_b200.models   <-> deepclr.models      (build_model, load_trained_model, ModelInferenceHelper, ...)
  deepclr_b200.config   <-> deepclr.config      (Config, load_model_config, ...)
  deepclr_b200.pointnet2 <-> pointnet2          (extern/Pointnet2.PyTorch as patched by extern/pointnet2.patch)
  deepclr_b200.torch_cluster <-> torch_cluster  (knn only)
All device work goes through the C ABI of libdeepclr_b200.so (include/deepclr_b200.h).
"""
import sys

__version__ = '0.1.0'


def install_as_deepclr() -> None:
    """Register this package under the reference's import names so `import deepclr.models`,
    `from pointnet2 import PointnetSAModuleMSG` and `from torch_cluster import knn` resolve to it."""
    import importlib
    for alias, target in (('pointnet2', 'deepclr_b200.pointnet2'),
                          ('pointnet2.pointnet2_utils', 'deepclr_b200.pointnet2.pointnet2_utils'),
                          ('pointnet2.pointnet2_modules', 'deepclr_b200.pointnet2.pointnet2_modules'),
                          ('torch_cluster', 'deepclr_b200.torch_cluster'),
                          ('deepclr', 'deepclr_b200'),
                          ('deepclr.config', 'deepclr_b200.config'),
                          ('deepclr.data', 'deepclr_b200.data'),
                          ('deepclr.data.labels', 'deepclr_b200.data.labels'),
                          ('deepclr.data.datasets', 'deepclr_b200.data.datasets'),
                          ('deepclr.data.datasets.kitti', 'deepclr_b200.data.datasets.kitti'),
                          ('deepclr.evaluation', 'deepclr_b200.evaluation'),
                          ('deepclr.evaluation.data', 'deepclr_b200.evaluation.data'),
                          ('deepclr.utils', 'deepclr_b200.utils'),
                          ('deepclr.models', 'deepclr_b200.models'),
                          ('deepclr.models.deepclr', 'deepclr_b200.models.deepclr'),
                          ('deepclr.models.base', 'deepclr_b200.models.base'),
                          ('deepclr.models.helper', 'deepclr_b200.models.helper')):
        sys.modules.setdefault(alias, importlib.import_module(target))


FPS_TIE_RULES = {'lowest_index': 0, 'reference': 1}
_fps_tie_rule = 'reference'


def set_fps_tie_rule(rule: str) -> None:
    """Host-side default of how farthest point sampling orders bit-equal maxima: what `furthest_point_sample` and
    models without an own `fps_tie_rule` pass down (the C ABI itself holds no global: the rule is an argument of
    dclr_fps and a property of a model handle, include/deepclr_b200.h).
    'reference' (default): the order of the reference's furthest_point_sampling_kernel (sampling_gpu.cu:113-203) --
    indices equal that kernel's on every input.  'lowest_index': the rule BASELINE.json states."""
    global _fps_tie_rule
    if rule not in FPS_TIE_RULES:
        raise ValueError(f'rule must be one of {sorted(FPS_TIE_RULES)}')
    _fps_tie_rule = rule


def get_fps_tie_rule() -> str:
    return _fps_tie_rule
