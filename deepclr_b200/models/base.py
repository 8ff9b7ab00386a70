"""BaseModel and ModelInferenceHelper: the inference API of the reference (deepclr/models/base.py:9-136)."""
import abc
import warnings
from typing import Dict, Optional, Tuple

import torch
import torch.nn as nn


class BaseModel(nn.Module, metaclass=abc.ABCMeta):
    """Interface of all registration models."""

    @abc.abstractmethod
    def get_input_dim(self) -> int:
        raise NotImplementedError

    @abc.abstractmethod
    def has_loss(self) -> bool:
        raise NotImplementedError

    @abc.abstractmethod
    def get_loss_weights(self) -> Dict:
        raise NotImplementedError

    @abc.abstractmethod
    def forward(self, x: torch.Tensor, is_feat: bool = False, m: Optional[torch.Tensor] = None,
                y: Optional[torch.Tensor] = None, debug: bool = False)\
            -> Tuple[torch.Tensor, Optional[torch.Tensor], Optional[Dict]]:
        """x: cloud batch [2B,N,C] (templates first) or features if is_feat; returns (y_pred, loss, debug)."""
        raise NotImplementedError

    @abc.abstractmethod
    def cloud_features(self, x: torch.Tensor, m: Optional[torch.Tensor] = None) -> torch.Tensor:
        raise NotImplementedError


class ModelInferenceHelper:
    """Sequential (odometry) and pairwise prediction with the reference's semantics: extra point channels are
    truncated with a warning, too few raise; sequential mode caches the previous frame's set-abstraction
    features; clouds of different size are randomly sub-sampled to the smaller one."""

    def __init__(self, model: BaseModel, is_sequential: bool = False, use_cuda_graph: bool = False):
        """`use_cuda_graph` (not in the reference): predictions of CUDA clouds replay the work from a CUDA graph captured
        per input shape -- pairwise: the whole forward (GraphedForward); sequential: the per-frame step = set abstraction
        of the new scan + merge/output against the cached features + cache update (GraphedSequentialStep) -- for streams
        of equally sized clouds, e.g. after the device-side subsample."""
        self._model = model
        self._input_dim = model.get_input_dim()
        self._model.eval()
        self._is_sequential = is_sequential
        self._state: Optional[torch.Tensor] = None
        self._use_graph = bool(use_cuda_graph)
        self._graphs: Dict = {}
        self._seq_graph = None                       # GraphedSequentialStep of the current cloud shape

    def has_state(self) -> bool:
        return self._state is not None or (self._seq_graph is not None and self._seq_graph._has_prev)

    def reset_state(self) -> None:
        self._state = None
        if self._seq_graph is not None:
            self._seq_graph.reset()

    def _fit_dim(self, cloud: torch.Tensor, what: str) -> torch.Tensor:
        if cloud.shape[1] > self._input_dim:
            warnings.warn(f"Truncate {what} point cloud from dimension {cloud.shape[1]} "
                          f"to required dimension {self._input_dim}.")
            return cloud[:, :self._input_dim]
        if cloud.shape[1] < self._input_dim:
            raise RuntimeError(f"Wrong point dimension in {what}.")
        return cloud

    def predict(self, source: torch.Tensor, template: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
        source = self._fit_dim(source, 'source')
        if template is not None:
            template = self._fit_dim(template, 'template')
        with torch.no_grad():
            if self._is_sequential:
                if template is not None:
                    raise RuntimeError("Only the source cloud is required for sequential prediction.")
                if self._use_graph and source.is_cuda:
                    from .stream import GraphedSequentialStep
                    sg = self._seq_graph
                    if sg is None or tuple(sg.static_x.shape[1:]) != tuple(source.shape):
                        carry = None
                        if sg is not None and sg._has_prev:       # a new cloud size mid-sequence: keep the cached features
                            carry = sg._prev
                        sg = self._seq_graph = GraphedSequentialStep(self._model, source)
                        if carry is not None:
                            sg._prev, sg._has_prev = carry.clone(), True
                    y = sg.step(source)
                    return None if y is None else y[0, :].clone()
                feat = self._model.cloud_features(source.unsqueeze(0))[0, ...]
                previous, self._state = self._state, feat
                if previous is None:
                    return None
                y, _, _ = self._model.forward(self.stack(previous, feat), is_feat=True)
                return y[0, :]
            if template is None:
                raise RuntimeError("Source and template clouds are required for non-sequential prediction.")
            x = self.stack(template, source)
            if self._use_graph and x.is_cuda:
                key = tuple(x.shape)
                if key not in self._graphs:
                    from .stream import GraphedForward
                    if len(self._graphs) >= 4:                       # a handful of shapes at most: drop the oldest
                        self._graphs.pop(next(iter(self._graphs)))
                    self._graphs[key] = GraphedForward(self._model, x)
                return self._graphs[key](x)[0, :].clone()
            y, _, _ = self._model.forward(x, is_feat=False)
            return y[0, :]

    @staticmethod
    def stack(template: torch.Tensor, source: torch.Tensor) -> torch.Tensor:
        """[2, N, C] batch, template first; the larger cloud is randomly reduced to the smaller size."""
        n = min(template.shape[0], source.shape[0])
        if template.is_cuda and source.is_cuda and template.shape[0] != source.shape[0]:
            # device-resident clouds: one native subsample kernel writes the [2, n, C] batch (no host permutation,
            # no index upload); the smaller cloud keeps its order as in the reference
            from ..data import device as dev_ops
            lengths = [int(template.shape[0]), int(source.shape[0])]
            rc = dev_ops.RaggedClouds(torch.cat((template, source), 0).float(),
                                      torch.tensor([0, lengths[0], lengths[0] + lengths[1]], dtype=torch.int32,
                                                   device=template.device), lengths)
            return dev_ops.subsample_clouds(rc, n, seed=int(torch.randint(0, 2 ** 62, (1,)).item()))
        if source.shape[0] > n:
            source = source[torch.randperm(source.shape[0])[:n], :]
        elif template.shape[0] > n:
            template = template[torch.randperm(template.shape[0])[:n], :]
        return torch.stack((template, source), 0)
