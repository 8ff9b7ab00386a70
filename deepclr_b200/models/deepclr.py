"""DeepCLR with the module tree, constructor arguments, state-dict keys and forward API of the reference
(deepclr/models/deepclr.py).  Two execution routes:

  * fused   : inference (grad disabled) on CUDA -> one C-ABI call (`dclr_forward_pairs` & co., include/deepclr_b200.h)
              that runs FPS, ball query, grouped MLP + max-pool, kNN, motion embedding and the pose head as
              hand-written sm_100a kernels.
  * composed: anything else the config API can express (training mode, augmentation matrix `m`, 2-level set
              abstraction, k = 0, append_features = False) -> the reference's op-by-op composition, with the native ops
              coming from the same C ABI through the `pointnet2` / `torch_cluster` drop-ins and the dense layers from
              torch.  There is no CPU route.
"""
import abc
import os
from typing import Any, Dict, List, Optional, Tuple, Union

import numpy as np
import torch
import torch.nn as nn

from .. import _lib
from ..config.config import Config
from ..data.labels import LabelType
from ..pointnet2 import PointnetSAModuleMSG
from ..torch_cluster import knn
from .base import BaseModel
from .compat_losses import (AccumulatedLoss, DeepCLRLoss, DeepCLRModule, TransformLoss,  # noqa: F401
                            TransformLossCalculation, TransformUncertaintyLoss, make_loss_layer, make_module, split_output)
from .helper import Conv1dMultiLayer, LinearMultiLayer
from .native import NativeModel


def split_features(x: torch.Tensor) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """[B,C,N] -> xyz [B,N,3] contiguous, features [B,C-3,N] or None."""
    xyz = x[:, :3, :].transpose(1, 2).contiguous()
    features = x[:, 3:, :].contiguous() if x.size(1) > 3 else None
    return xyz, features


def merge_features(xyz: torch.Tensor, features: Optional[torch.Tensor]) -> torch.Tensor:
    """xyz [B,n,3], features [B,F,n] -> [B,3+F,n] (xyz rows first)."""
    if features is None:
        return xyz.transpose(1, 2)
    return torch.cat((xyz.transpose(1, 2), features), dim=1)


class SetAbstraction(DeepCLRModule):
    """PointNet++ multi-scale set abstraction of every cloud."""

    def __init__(self, input_dim: int, point_dim: int, mlps: List[List[List[int]]], npoint: List[int],
                 radii: List[List[float]], nsamples: List[List[int]], batch_norm: bool = False, **_kwargs: Any):
        super().__init__()
        assert point_dim == 3
        assert len(mlps) == len(npoint) == len(radii) == len(nsamples)
        assert 0 < len(mlps) <= 2
        self._point_dim = point_dim
        self._output_feat_dim = int(np.sum([x[-1] for x in mlps[-1]]))
        self._spec = dict(npoint=list(npoint), radii=[list(r) for r in radii], nsamples=[list(n) for n in nsamples],
                          mlps=[[list(m) for m in lvl] for lvl in mlps], input_dim=input_dim)
        self._sa0 = PointnetSAModuleMSG(npoint=npoint[0], radii=radii[0], nsamples=nsamples[0],
                                        mlps=[[input_dim - point_dim, *x] for x in mlps[0]], use_xyz=True,
                                        bn=batch_norm)
        if len(npoint) == 2:
            self._sa1 = PointnetSAModuleMSG(npoint=npoint[1], radii=radii[1], nsamples=nsamples[1],
                                            mlps=[[*x] for x in mlps[1]], use_xyz=True, bn=batch_norm)
        else:
            self._sa1 = None

    def output_dim(self) -> int:
        return 3 + self._output_feat_dim

    def forward(self, clouds: torch.Tensor, *_args: Any) -> torch.Tensor:
        xyz, features = split_features(clouds)
        xyz, features = self._sa0(xyz, features)
        if self._sa1 is not None:
            xyz, features = self._sa1(xyz, features)
        return merge_features(xyz, features)


class GroupingModule(abc.ABC, nn.Module):
    @abc.abstractmethod
    def forward(self, cloud0: torch.Tensor, cloud1: torch.Tensor)\
            -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor]:
        raise NotImplementedError


def _rows(cloud: torch.Tensor) -> torch.Tensor:
    return cloud.transpose(1, 2).contiguous().view(-1, cloud.shape[1])


class GlobalGrouping(GroupingModule):
    """Every template point grouped with all source points of its pair."""

    def forward(self, cloud0, cloud1):
        pts0, pts1 = _rows(cloud0), _rows(cloud1)
        B, n0, n1 = cloud0.shape[0], cloud0.shape[2], cloud1.shape[2]
        idx0 = torch.arange(B * n0, device=pts0.device).view(-1, 1).expand(-1, n1)
        idx1 = torch.arange(B * n1, device=pts1.device).view(B, 1, n1).expand(-1, n0, -1).reshape(B * n0, n1)
        return pts0, pts1, pts0[idx0], pts1[idx1]


class KnnGrouping(GroupingModule):
    """Every template point grouped with its k nearest source points."""

    def __init__(self, point_dim: int, k: int):
        super().__init__()
        self._point_dim = point_dim
        self._k = k

    def forward(self, cloud0, cloud1):
        pts0, pts1 = _rows(cloud0), _rows(cloud1)
        batch0 = torch.arange(cloud0.shape[0], device=pts0.device).repeat_interleave(cloud0.shape[2])
        batch1 = torch.arange(cloud1.shape[0], device=pts1.device).repeat_interleave(cloud1.shape[2])
        group_index = knn(pts1[:, :self._point_dim].contiguous().detach(),
                          pts0[:, :self._point_dim].contiguous().detach(),
                          k=self._k, batch_x=batch1, batch_y=batch0).view(2, pts0.shape[0], self._k)
        return pts0, pts1, pts0[group_index[0]], pts1[group_index[1]]


class MotionEmbeddingBase(nn.Module):
    """Flow embedding between template (clouds0) and source (clouds1) features."""

    def __init__(self, input_dim: int, point_dim: int, k: int, radius: float, mlp: List[int],
                 append_features: bool = True, batch_norm: bool = False, **_kwargs: Any):
        super().__init__()
        self._point_dim = point_dim
        self._append_features = append_features
        self._k = k
        self._grouping = GlobalGrouping() if k == 0 else KnnGrouping(point_dim, k)
        first = point_dim + 2 * (input_dim - point_dim) if append_features else input_dim
        self._conv = Conv1dMultiLayer([first, *mlp], batch_norm=batch_norm)
        self._radius = radius

    def output_dim(self) -> int:
        return self._point_dim + self._conv.output_dim()

    def forward(self, clouds0: torch.Tensor, clouds1: torch.Tensor) -> torch.Tensor:
        d = self._point_dim
        pts0, pts1, g0, g1 = self._grouping(clouds0, clouds1)
        pos_diff = g1[:, :, :d] - g0[:, :, :d]
        if self._append_features:
            merged = torch.cat((pos_diff, g0[:, :, d:], g1[:, :, d:]), dim=2)
        else:
            merged = torch.cat((pos_diff, g1[:, :, d:] - g0[:, :, d:]), dim=2)
        feat = self._conv(merged.transpose(1, 2))
        if self._radius > 0.0:
            outside = (torch.norm(pos_diff, dim=2) >= self._radius).unsqueeze(1)
            feat = feat.masked_fill(outside, 0.0)
        feat, _ = torch.max(feat, dim=2)
        out = torch.cat((pts0[:, :d], feat), dim=1)
        return out.view(clouds0.shape[0], -1, out.shape[1]).transpose(1, 2).contiguous()


class MotionEmbedding(DeepCLRModule):
    """Motion embedding of a batch sorted [template1, ..., templateB, source1, ..., sourceB]."""

    def __init__(self, **kwargs: Any):
        super().__init__()
        self._embedding = MotionEmbeddingBase(**kwargs)

    def output_dim(self):
        return self._embedding.output_dim()

    def forward(self, clouds: torch.Tensor) -> torch.Tensor:
        half = int(clouds.shape[0] / 2)
        return self._embedding(clouds[:half, ...], clouds[half:, ...])


class OutputSimple(DeepCLRModule):
    """Mini-PointNet over the embedded points + fully connected regression of the pose label."""

    def __init__(self, input_dim: int, label_type: LabelType, mlp: List[int], linear: List[int],
                 batch_norm: bool = False, dropout: bool = False, **_kwargs: Any):
        super().__init__()
        self._label_type = label_type
        self.conv = Conv1dMultiLayer([input_dim, *mlp], batch_norm=batch_norm)
        self.linear = LinearMultiLayer(linear, batch_norm=batch_norm, dropout_keep=dropout, dropout_last=True)
        self.output = nn.Linear(linear[-1], label_type.dim, bias=True)
        nn.init.xavier_uniform_(self.output.weight)
        if label_type.bias is not None:
            with torch.no_grad():
                self.output.bias.copy_(torch.tensor(label_type.bias))

    def output_dim(self) -> int:
        return self._label_type.dim

    def _output_activation(self, x: torch.Tensor) -> torch.Tensor:
        if self._label_type == LabelType.POSE3D_QUAT:
            return torch.cat((x[:, :3], torch.sigmoid(x[:, 3:4]), torch.tanh(x[:, 4:])), dim=1)
        if self._label_type == LabelType.POSE3D_DUAL_QUAT:
            return torch.cat((torch.sigmoid(x[:, 0:1]), torch.tanh(x[:, 1:4]), x[:, 4:]), dim=1)
        return x

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x, _ = torch.max(self.conv(x), dim=2)
        return self._output_activation(self.output(self.linear(x)))


def _as_cfg(c: Union[Config, dict, None]) -> Optional[Config]:
    if c is None or isinstance(c, Config):
        return c
    return Config._from_dict(c)


class DeepCLR(BaseModel):
    """Main network: cloud features -> motion embedding -> pose regression."""
    _loss_layer: Optional[DeepCLRLoss]

    def __init__(self, input_dim: int, label_type: LabelType, cloud_features: Config, merge: Config, output: Config,
                 transform: Optional[Config] = None, loss: Optional[Config] = None, **kwargs: Any):
        super().__init__()
        self._input_dim = input_dim
        self._label_type = label_type
        cloud_features, merge, output, transform = map(_as_cfg, (cloud_features, merge, output, transform))
        transform_layer = None if transform is None else make_module(DeepCLRModule, transform, input_dim=input_dim, **kwargs)
        feat_in = input_dim if transform_layer is None else transform_layer.output_dim()
        cloud_feat_layer = make_module(DeepCLRModule, cloud_features, input_dim=feat_in, **kwargs)
        merge_layer = make_module(DeepCLRModule, merge, input_dim=cloud_feat_layer.output_dim(), **kwargs)
        output_layer = make_module(DeepCLRModule, output, input_dim=merge_layer.output_dim(), label_type=label_type,
                                   **kwargs)
        self._cloud_layers = nn.Sequential(cloud_feat_layer) if transform_layer is None \
            else nn.Sequential(transform_layer, cloud_feat_layer)
        self._merge_layers = nn.Sequential(merge_layer, output_layer)
        self._loss_layer = make_loss_layer(loss, label_type, _as_cfg, **kwargs)
        # fused route state
        # numerics of the fused route: 'fp16' (default) | 'bf16' = tcgen05 tensor cores, 16-bit operands, fp32 accumulation;
        # 'fp32' = SIMT fp32 everywhere.  Why fp16 is the default: its 11-bit significand is TF32's, which is what the
        # reference's own convolutions run in on any GPU since Ampere (torch.backends.cudnn.allow_tf32 defaults to True);
        # measured pose deviation from strict fp32: reference with TF32 0.069 deg max, this fp16 mode 0.093 deg max
        # (profiles/r2_precision_class.json).  'fp32' is the bit-faithful, ~12x slower choice.
        self.numerics = os.environ.get('DCLR_NUMERICS', 'fp16')
        self.fps_tie_rule: Optional[str] = None    # None = deepclr_b200.get_fps_tie_rule() at call time ('reference')
        self._native: Optional[NativeModel] = None
        self._native_key = None

    # ---- reference API
    def get_input_dim(self) -> int:
        return self._input_dim

    def has_loss(self) -> bool:
        return self._loss_layer is not None

    def get_loss_weights(self) -> Dict:
        return self._loss_layer.get_weights() if self._loss_layer is not None else {}

    def forward(self, x: torch.Tensor, is_feat: bool = False, m: Optional[torch.Tensor] = None,
                y: Optional[torch.Tensor] = None, debug: bool = False)\
            -> Tuple[torch.Tensor, Optional[torch.Tensor], Optional[Dict]]:
        if self._fused_ok(x, m):
            native = self._get_native(x.device)
            if is_feat:
                y_pred = native.forward_features(x)
            else:
                y_pred, x, _ = native.forward_pairs(x, want_feat=debug)
            model_aux: Dict = dict()
        else:
            if not is_feat:
                x = self.cloud_features(x, m=m)
            y_pred, model_aux = split_output(self._merge_layers(x))
        if self._loss_layer is not None and y is not None:
            loss, loss_aux = split_output(self._loss_layer(y_pred, y, **model_aux))
            debug_output = {**model_aux, **loss_aux, 'x_aug': x} if debug else None
        else:
            loss, debug_output = None, None
        return y_pred, loss, debug_output

    def cloud_features(self, x: torch.Tensor, m: Optional[torch.Tensor] = None) -> torch.Tensor:
        if self._fused_ok(x, m):
            return self._get_native(x.device).cloud_features(x)
        if m is not None:
            # tgm.transform_points(m, xyz) of the reference (deepclr.py:512-514): homogeneous 4x4 applied per cloud
            dim = m.shape[-1] - 1
            x[:, :, :dim] = x[:, :, :dim] @ m[:, :dim, :dim].transpose(1, 2) + m[:, :dim, dim].unsqueeze(1)
        return self._cloud_layers(x.transpose(1, 2))

    # ---- fused route
    def replica(self) -> 'DeepCLR':
        """An independent copy of the model (own parameters, own native handle and workspace): forwards of the original
        and of a replica may overlap on different CUDA streams (one handle serialises its calls: include/deepclr_b200.h,
        threading contract).  Used by HostPipeline(replicas=2) so that the latency-bound sampling of batch i+1 runs
        beside the dense stages of batch i."""
        import copy
        native, key = self._native, self._native_key
        self._native, self._native_key = None, None
        try:
            twin = copy.deepcopy(self)
        finally:
            self._native, self._native_key = native, key
        return twin

    def graph_token(self, device: torch.device):
        """Identity of everything a CUDA graph captured from this model's fused forward points at: the native handle
        (rebuilt when parameters, numerics or the tie rule change) and its workspace generation (bumped when a larger
        shape re-allocates it).  GraphedForward / HostPipeline / ModelInferenceHelper re-capture when it changes."""
        native = self._get_native(device)
        return (id(native), native.generation())

    def _fused_supported(self) -> bool:
        sa, me, out = self._cloud_layers[-1], self._merge_layers[0], self._merge_layers[1]
        if len(self._cloud_layers) != 1 or not isinstance(sa, SetAbstraction) or sa._sa1 is not None:
            return False
        if not isinstance(me, MotionEmbedding) or not isinstance(out, OutputSimple):
            return False
        emb = me._embedding
        spec = sa._spec
        if not (emb._append_features and len(spec['radii'][0]) in (1, 2)
                and len(emb._conv._sequential) == 3 and len(out.conv._sequential) == 5
                and len(out.linear._sequential) == 2):
            return False
        # every constraint of the native kernels (csrc/sa_mlp.cu: sa_mlp_launch, csrc/model.cu: dclr_model_create,
        # csrc/mlp_tc.cu: tc_create); any other legal reference config runs the composed route
        if spec['input_dim'] not in (3, 4) or any(list(mm) != [16, 16, 32] for mm in spec['mlps'][0]):
            return False
        if not 1 <= emb._k <= min(128, spec['npoint'][0]):
            return False
        if self.numerics != 'fp32':
            F = 32 * len(spec['mlps'][0])
            merge = [b.output_dim() for b in emb._conv._sequential]
            head = [b.output_dim() for b in out.conv._sequential]
            if merge[:2] != [128, 128] or merge[2] not in (128, 256):
                return False
            if emb._conv._sequential[0]._sequential[0].in_channels != 2 * F + 3:
                return False
            if any(h % 128 for h in head) or any(h > 512 for h in head[:-1]):
                return False
        return True

    def _fused_ok(self, x: torch.Tensor, m: Optional[torch.Tensor]) -> bool:
        if not x.is_cuda:
            raise RuntimeError('deepclr_b200 has no CPU path: move the model and its inputs to a CUDA device')
        if torch.is_grad_enabled() or m is not None:
            return False
        if self.training and any(isinstance(mod, (nn.BatchNorm1d, nn.BatchNorm2d, nn.Dropout))
                                 for mod in self.modules()):
            return False
        return self._fused_supported()

    def _dense_layers(self):
        """(layer, batch-norm) pairs in state-dict order."""
        sa, me, out = self._cloud_layers[-1], self._merge_layers[0], self._merge_layers[1]
        pairs = []
        for mlp in sa._sa0.mlps:
            for block in mlp:
                pairs.append((block.conv, block.bn.bn if hasattr(block, 'bn') else None))
        for stack in (me._embedding._conv, out.conv, out.linear):
            for block in stack._sequential:
                seq = block._sequential
                pairs.append((seq[0], seq[1] if len(seq) > 1 else None))
        pairs.append((out.output, None))
        return pairs

    def _get_native(self, device: torch.device) -> NativeModel:
        pairs = self._dense_layers()
        import deepclr_b200
        tie_rule = self.fps_tie_rule or deepclr_b200.get_fps_tie_rule()
        key = (str(device), self.numerics, tie_rule) + tuple((p.data_ptr(), p._version) for layer, _ in pairs
                                                   for p in layer.parameters())
        if self._native is None or key != self._native_key:
            sa, me, out = self._cloud_layers[-1], self._merge_layers[0], self._merge_layers[1]
            spec, emb = sa._spec, me._embedding
            arch = _lib.DclrArch()
            arch.input_dim = spec['input_dim']
            arch.npoint = spec['npoint'][0]
            arch.num_scales = len(spec['radii'][0])
            for s in range(arch.num_scales):
                arch.radii[s] = spec['radii'][0][s]
                arch.nsamples[s] = spec['nsamples'][0][s]
                for l in range(3):
                    arch.sa_mlp[s][l] = spec['mlps'][0][s][l]
            arch.knn_k = emb._k
            arch.merge_radius = float(emb._radius)
            for l, block in enumerate(emb._conv._sequential):
                arch.merge_mlp[l] = block.output_dim()
            for l, block in enumerate(out.conv._sequential):
                arch.out_mlp[l] = block.output_dim()
            arch.out_linear[0] = out.linear._sequential[0]._sequential[0].in_features
            for l, block in enumerate(out.linear._sequential):
                arch.out_linear[l + 1] = block.output_dim()
            arch.label_type = self._label_type.native_id
            self._native = NativeModel(arch, pairs, device, self.numerics, tie_rule)
            self._native_key = key
        return self._native
