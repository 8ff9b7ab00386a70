"""In-model loss modules of the reference config API (`loss:` block of a model config, deepclr/models/deepclr.py:297-409).

COMPATIBILITY ONLY: training is outside the accelerated path; these classes exist so that a
reference config with a `loss:` entry builds, `forward(x, y=...)` returns a loss and a reference checkpoint with learnt
loss weights (`_loss_layer._sx`, `_loss_layer._sq`) loads.  What is fixed by that contract: the class names (configs name
them), the constructor keywords, the parameter names `_sx` / `_sq`, `get_weights()`.  Everything else is this file's own.
"""
import abc
from typing import Any, Callable, Dict, Iterable, Optional, Tuple

import torch
import torch.nn as nn

from ..data.labels import LabelType
from ..utils.factory import factory
from ..utils.metrics import rot_loss, trans_loss


class DeepCLRModule(nn.Module, metaclass=abc.ABCMeta):
    """Base class of the configurable sub-modules (looked up by class name)."""

    @abc.abstractmethod
    def output_dim(self) -> int:
        raise NotImplementedError


def split_output(output: Any) -> Tuple[Any, Dict]:
    """A module may return `value` or `(value, aux_dict)`."""
    if not isinstance(output, (list, tuple)):
        return output, {}
    value, aux = output
    return value, aux


def make_module(base: type, cfg, *args: Any, **kwargs: Any):
    """Instantiate the subclass of `base` that a config node names, with the node's params."""
    return factory(base, cfg.name, *args, **cfg.params, **kwargs)


class TransformLossCalculation(nn.Module):
    """Per-sample translation and rotation label errors, optionally averaged over the batch; non-finite values raise."""

    def __init__(self, label_type: LabelType, p: int, reduction: Optional[str] = 'mean'):
        super().__init__()
        self._label_type, self._p, self._reduction = label_type, p, reduction

    def forward(self, y_pred: torch.Tensor, y: torch.Tensor):
        terms = {'translation': trans_loss(y_pred, y, self._label_type, p=self._p, reduction='none'),
                 'rotation': rot_loss(y_pred, y, self._label_type, p=self._p, reduction='none')}
        if self._reduction == 'mean':
            terms = {k: v.mean(dim=0) for k, v in terms.items()}
        for what, v in terms.items():
            if not bool(torch.isfinite(v).all()):
                raise RuntimeError(f"TransformLoss: {what} loss is nan or inf:\ny_pred = \n{y_pred}\ny = \n{y}")
        if self._reduction == 'mean':
            return terms['translation'][0], terms['rotation'][0]
        return torch.cat((terms['translation'], terms['rotation']), dim=1)


class DeepCLRLoss(DeepCLRModule, metaclass=abc.ABCMeta):
    def output_dim(self) -> int:
        return 1

    @abc.abstractmethod
    def get_weights(self) -> Dict:
        raise NotImplementedError


class _WeightedPoseLoss(DeepCLRLoss, metaclass=abc.ABCMeta):
    """translation and rotation error combined by `_combine`; subclasses choose fixed or learnt weights."""

    def __init__(self, label_type: LabelType, p: int, **_unused: Any):
        super().__init__()
        self._transform_loss = TransformLossCalculation(label_type, p, reduction='mean')

    @abc.abstractmethod
    def _combine(self, t: torch.Tensor, r: torch.Tensor) -> torch.Tensor:
        raise NotImplementedError

    def forward(self, y_pred: torch.Tensor, y: torch.Tensor, **_aux: Any) -> torch.Tensor:
        return self._combine(*self._transform_loss(y_pred, y))


class TransformLoss(_WeightedPoseLoss):
    """sx * translation + sq * rotation with constant weights."""

    def __init__(self, label_type: LabelType, p: int, sx: float, sq: float, **kwargs: Any):
        super().__init__(label_type, p, **kwargs)
        self._sx, self._sq = sx, sq

    def get_weights(self) -> Dict:
        return {}

    def _combine(self, t, r):
        return self._sx * t + self._sq * r


class TransformUncertaintyLoss(_WeightedPoseLoss):
    """Homoscedastic-uncertainty weighting: exp(-s) * loss + s per term, s learnt (parameters `_sx`, `_sq`)."""

    def __init__(self, label_type: LabelType, p: int, sx: float, sq: float, **kwargs: Any):
        super().__init__(label_type, p, **kwargs)
        self._sx = nn.Parameter(torch.tensor([float(sx)]))
        self._sq = nn.Parameter(torch.tensor([float(sq)]))

    def get_weights(self) -> Dict:
        return {'sx': self._sx.item(), 'sq': self._sq.item()}

    def _combine(self, t, r):
        return sum(torch.exp(-s) * term + s for s, term in ((self._sx, t), (self._sq, r)))


class AccumulatedLoss(DeepCLRLoss):
    """Sum of several loss modules (a `loss:` list in the config); their weight names must not clash."""

    def __init__(self, modules: Iterable[nn.Module]):
        super().__init__()
        self.loss_list = nn.ModuleList(modules)

    def get_weights(self) -> Dict:
        merged: Dict = {}
        for part in self.loss_list:
            w = part.get_weights()
            if merged.keys() & w.keys():
                raise RuntimeError("Duplicate loss keys")
            merged.update(w)
        return merged

    def forward(self, *args: Any, **kwargs: Any) -> torch.Tensor:
        return torch.stack([part(*args, **kwargs) for part in self.loss_list]).sum()


def make_loss_layer(loss_cfg, label_type: LabelType, as_cfg: Callable, **kwargs: Any) -> Optional[DeepCLRLoss]:
    """None, one loss config or a list of them -> the model's loss layer."""
    if loss_cfg is None:
        return None
    if isinstance(loss_cfg, list):
        return AccumulatedLoss([make_module(DeepCLRLoss, as_cfg(c), label_type=label_type, **kwargs) for c in loss_cfg])
    return make_module(DeepCLRLoss, as_cfg(loss_cfg), label_type=label_type, **kwargs)
