"""Import the reference's own Python model code from /root/reference on CPU.  TEST INFRASTRUCTURE ONLY.

Works only where /root/reference is mounted (the build container).  Third-party packages the reference imports
but this image lacks are stubbed in sys.modules; the five native ops the reference calls are bound, at function
level, to the CPU restatements in oracle/ops.py.  Everything else (module classes, layer order, init, concat
order, masks, activations) is the reference's unmodified code, which is what makes the goldens authoritative for
the dense part of the path.
"""
import importlib
import os
import sys
import types

import numpy as np
import torch

from . import ops

REF_ROOT = os.environ.get('DCLR_REFERENCE_ROOT', '/root/reference')

_STUBS = ['torch_cluster', 'torchgeometry', 'pointnet2_cuda', 'transforms3d', 'dataflow', 'dataflow.dataflow',
          'ignite', 'ignite._utils', 'ignite.engine', 'ignite.handlers', 'ignite.metrics', 'ranger',
          'ranger.ranger2020', 'tensorboardX', 'lmdb', 'pykitti', 'open3d']


class _Anything(types.ModuleType):
    """Module stub: any attribute resolves to a dummy class so `from x import Y` / subclassing succeed."""
    def __getattr__(self, name):
        if name.startswith('__'):
            raise AttributeError(name)
        obj = type(name, (), {'__init__': lambda self, *a, **k: None})
        setattr(self, name, obj)
        return obj


def available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, 'deepclr', 'models'))


def _np(t):
    return t.detach().cpu().numpy()


class _Fn:
    """Function-level shims with the reference signatures (pointnet2_utils.py:36,73,197,228; deepclr.py:164)."""
    fps_mode = 'reference'   # 'reference' = block-tree tie-break (sampling_gpu.cu), 'lowidx' = stated rule

    @staticmethod
    def furthest_point_sample(xyz, npoint):
        f = ops.fps if _Fn.fps_mode == 'reference' else ops.fps_lowidx
        return torch.from_numpy(f(_np(xyz), npoint))

    @staticmethod
    def gather_operation(features, idx):
        return torch.from_numpy(ops.gather(_np(features), _np(idx)))

    @staticmethod
    def ball_query(radius, nsample, xyz, new_xyz):
        return torch.from_numpy(ops.ball_query(radius, nsample, _np(xyz), _np(new_xyz)))

    @staticmethod
    def grouping_operation(features, idx):
        return torch.from_numpy(ops.group(_np(features), _np(idx)))

    @staticmethod
    def knn(x, y, k, batch_x=None, batch_y=None, cosine=False, num_workers=1):
        B = 1 if batch_x is None else int(batch_x.max().item()) + 1
        return torch.from_numpy(ops.knn(_np(x), _np(y), k, B))


_loaded = None


def load():
    """Returns a namespace with the reference's DeepCLR, LabelType, Config classes (CPU)."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError(f'reference tree not found at {REF_ROOT}')
    for name in _STUBS:
        if name not in sys.modules:
            sys.modules[name] = _Anything(name)
    sys.modules['torch_cluster'].knn = _Fn.knn
    for p in (os.path.join(REF_ROOT, 'extern', 'Pointnet2.PyTorch'), REF_ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    pu = importlib.import_module('pointnet2.pointnet2_utils')
    pu.furthest_point_sample = _Fn.furthest_point_sample
    pu.gather_operation = _Fn.gather_operation
    pu.ball_query = _Fn.ball_query
    pu.grouping_operation = _Fn.grouping_operation
    pm = importlib.import_module('pointnet2.pointnet2_modules')
    # extern/pointnet2.patch:7-9 makes `from pointnet2 import PointnetSAModuleMSG` work; emulate the patched __init__
    sys.modules['pointnet2'].PointnetSAModuleMSG = pm.PointnetSAModuleMSG
    importlib.import_module('deepclr.config')      # must come first: deepclr.models <-> deepclr.config is circular
    dc = importlib.import_module('deepclr.models.deepclr')
    labels = importlib.import_module('deepclr.data.labels')
    cfgmod = importlib.import_module('deepclr.config.config')
    _loaded = types.SimpleNamespace(DeepCLR=dc.DeepCLR, deepclr_module=dc, LabelType=labels.LabelType,
                                    Config=cfgmod.Config, Fn=_Fn)
    return _loaded


def build_reference_model(cfg: dict, seed: int = 0):
    """Instantiate the reference DeepCLR from a model_config dict with its own initialisers, seeded."""
    ref = load()

    def to_cfg(d):
        c = ref.Config(allow_dynamic_params=True)
        for k, v in d.items():
            setattr(c, k, to_cfg(v) if isinstance(v, dict) else v)
        return c

    params = {k: (to_cfg(v) if isinstance(v, dict) else v) for k, v in cfg['params'].items()}
    torch.manual_seed(seed)
    label = ref.LabelType[str(cfg['label_type']).upper()]
    model = ref.DeepCLR(input_dim=cfg['input_dim'], point_dim=cfg['point_dim'], label_type=label, **params)
    return model.eval()
