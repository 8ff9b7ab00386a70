"""Native (C-ABI) model handle for a DeepCLR module tree: packs the parameters in state-dict order, creates the
`dclr_model` and keeps it in sync with the torch parameters."""
import ctypes
from typing import List, Optional, Tuple

import torch
import torch.nn as nn

from .. import _lib


def _fold(layer: nn.Module, norm: Optional[nn.Module]) -> Tuple[torch.Tensor, torch.Tensor]:
    """(W [out,in], b [out]) of a 1x1 conv / linear with an optional eval-mode BatchNorm folded in."""
    w = layer.weight.detach().float().reshape(layer.weight.shape[0], -1)
    b = layer.bias.detach().float() if layer.bias is not None else w.new_zeros(w.shape[0])
    if norm is not None:
        scale = norm.weight.detach().float() / torch.sqrt(norm.running_var.float() + norm.eps)
        w = w * scale[:, None]
        b = (b - norm.running_mean.float()) * scale + norm.bias.detach().float()
    return w, b


def _first_bn(seq) -> Optional[nn.Module]:
    for mod in seq.modules():
        if isinstance(mod, (nn.BatchNorm1d, nn.BatchNorm2d)):
            return mod
    return None


class NativeModel:
    """Owns one `dclr_model*` on one device."""

    def __init__(self, arch: _lib.DclrArch, layers: List[Tuple[nn.Module, Optional[nn.Module]]], device: torch.device,
                 numerics: str, fps_tie_rule: str = 'reference'):
        blobs = []
        for layer, norm in layers:
            w, b = _fold(layer, norm)
            blobs += [w.reshape(-1).cpu(), b.reshape(-1).cpu()]
        flat = torch.cat(blobs).contiguous()
        self.device = device
        self._handle = ctypes.c_void_p()
        with torch.cuda.device(device):
            _lib.check(_lib.lib().dclr_model_create(ctypes.byref(arch), ctypes.c_void_p(flat.data_ptr()),
                                                    flat.numel(), ctypes.byref(self._handle)), 'dclr_model_create')
            _lib.check(_lib.lib().dclr_model_set_numerics(self._handle, {'fp32': 0, 'bf16': 1, 'fp16': 2}[numerics]),
                       'dclr_model_set_numerics')
            from .. import FPS_TIE_RULES
            _lib.check(_lib.lib().dclr_model_set_fps_tie_rule(self._handle, FPS_TIE_RULES[fps_tie_rule]),
                       'dclr_model_set_fps_tie_rule')
        self.numerics = numerics
        self.fps_tie_rule = fps_tie_rule
        self.label_dim = {0: 6, 1: 7, 2: 8}[arch.label_type]
        self.npoint = arch.npoint
        self.feat_dim = 3 + sum(arch.sa_mlp[s][2] for s in range(arch.num_scales))
        self.motion_dim = 3 + arch.merge_mlp[2]
        self.knn_k = arch.knn_k
        self.nsamples = [arch.nsamples[s] for s in range(arch.num_scales)]

    def __del__(self):
        try:
            if self._handle:
                _lib.lib().dclr_model_destroy(self._handle)
                self._handle = ctypes.c_void_p()
        except Exception:
            pass

    def _prep(self, x: torch.Tensor) -> torch.Tensor:
        _lib.require_cuda(x)
        if x.device != self.device:
            raise RuntimeError(f'input on {x.device} but native model on {self.device}')
        return x.contiguous().float()

    def cloud_features(self, x: torch.Tensor) -> torch.Tensor:
        x = self._prep(x)
        G, N, _ = x.shape
        out = torch.empty((G, self.feat_dim, self.npoint), dtype=torch.float32, device=x.device)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().dclr_cloud_features(self._handle, _lib.ptr(x), G, N, _lib.ptr(out),
                                                      _lib.stream_ptr()), 'dclr_cloud_features')
        return out

    def forward_features(self, feat: torch.Tensor, want_motion: bool = False):
        feat = self._prep(feat)
        if feat.shape[0] % 2 or feat.shape[1] != self.feat_dim or feat.shape[2] != self.npoint:
            raise RuntimeError(f'features must be [2B,{self.feat_dim},{self.npoint}], got {tuple(feat.shape)}')
        B = feat.shape[0] // 2
        y = torch.empty((B, self.label_dim), dtype=torch.float32, device=feat.device)
        motion = torch.empty((B, self.motion_dim, self.npoint), dtype=torch.float32, device=feat.device) \
            if want_motion else None
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().dclr_forward_features(self._handle, _lib.ptr(feat), B, _lib.ptr(y),
                                                        _lib.ptr(motion), _lib.stream_ptr()), 'dclr_forward_features')
        return (y, motion) if want_motion else y

    def forward_pairs(self, x: torch.Tensor, want_feat: bool = False, want_motion: bool = False):
        x = self._prep(x)
        if x.shape[0] % 2:
            raise RuntimeError('batch must hold templates then sources: even number of clouds')
        B, N = x.shape[0] // 2, x.shape[1]
        y = torch.empty((B, self.label_dim), dtype=torch.float32, device=x.device)
        feat = torch.empty((2 * B, self.feat_dim, self.npoint), dtype=torch.float32, device=x.device) \
            if want_feat else None
        motion = torch.empty((B, self.motion_dim, self.npoint), dtype=torch.float32, device=x.device) \
            if want_motion else None
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().dclr_forward_pairs(self._handle, _lib.ptr(x), B, N, _lib.ptr(y), _lib.ptr(feat),
                                                     _lib.ptr(motion), _lib.stream_ptr()), 'dclr_forward_pairs')
        return y, feat, motion

    # ---- stage-level entries (include/deepclr_b200.h)
    def workspace_bytes(self, num_clouds: int, n_points: int) -> int:
        return int(_lib.lib().dclr_workspace_bytes(self._handle, num_clouds, n_points))

    def sa_msg(self, x: torch.Tensor, new_xyz: torch.Tensor) -> torch.Tensor:
        """x [G,N,C], centres new_xyz [G,npoint,3] -> features [G,F,npoint] (PointnetSAModuleMSG after its sampling)."""
        x, new_xyz = self._prep(x), self._prep(new_xyz)
        G, N, _ = x.shape
        if tuple(new_xyz.shape) != (G, self.npoint, 3):
            raise RuntimeError(f'new_xyz must be [{G},{self.npoint},3], got {tuple(new_xyz.shape)}')
        out = torch.empty((G, self.feat_dim - 3, self.npoint), dtype=torch.float32, device=x.device)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().dclr_sa_msg_fused(self._handle, _lib.ptr(x), _lib.ptr(new_xyz), G, N, _lib.ptr(out),
                                                    _lib.stream_ptr()), 'dclr_sa_msg_fused')
        return out

    def flow_embed(self, feat: torch.Tensor) -> torch.Tensor:
        """feat [2B,3+F,npoint] templates first -> motion embedding [B,3+C,npoint] (MotionEmbedding.forward)."""
        feat = self._prep(feat)
        if feat.shape[0] % 2 or feat.shape[1] != self.feat_dim or feat.shape[2] != self.npoint:
            raise RuntimeError(f'features must be [2B,{self.feat_dim},{self.npoint}], got {tuple(feat.shape)}')
        B = feat.shape[0] // 2
        out = torch.empty((B, self.motion_dim, self.npoint), dtype=torch.float32, device=feat.device)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().dclr_flow_embed_fused(self._handle, _lib.ptr(feat), B, _lib.ptr(out),
                                                        _lib.stream_ptr()), 'dclr_flow_embed_fused')
        return out

    def pose_head(self, motion: torch.Tensor) -> torch.Tensor:
        """motion [B,3+C,npoint] -> pose labels [B,label_dim] (OutputSimple.forward)."""
        motion = self._prep(motion)
        if motion.shape[1] != self.motion_dim or motion.shape[2] != self.npoint:
            raise RuntimeError(f'motion must be [B,{self.motion_dim},{self.npoint}], got {tuple(motion.shape)}')
        B = motion.shape[0]
        y = torch.empty((B, self.label_dim), dtype=torch.float32, device=motion.device)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().dclr_pose_head_fused(self._handle, _lib.ptr(motion), B, _lib.ptr(y),
                                                       _lib.stream_ptr()), 'dclr_pose_head_fused')
        return y

    def generation(self) -> int:
        """Workspace generation of the handle (include/deepclr_b200.h: dclr_model_generation)."""
        return int(_lib.lib().dclr_model_generation(self._handle))

    def last_launches(self) -> int:
        return int(_lib.lib().dclr_model_last_launches(self._handle))

    def set_timing(self, on: bool) -> None:
        _lib.check(_lib.lib().dclr_model_set_timing(self._handle, int(on)), 'dclr_model_set_timing')

    def collect_timing(self):
        """-> ({stage: summed ms}, calls) over the forwards recorded since the last collect (at most 64)."""
        n = _lib.lib().dclr_model_stage_count()
        buf = (ctypes.c_float * n)()
        calls = ctypes.c_int()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().dclr_model_collect_timing(self._handle, buf, ctypes.byref(calls)),
                       'dclr_model_collect_timing')
        return {_lib.lib().dclr_model_stage_name(i).decode(): float(buf[i]) for i in range(n)}, calls.value

    def debug_tensor(self, which: int, shape, dtype=torch.int32) -> torch.Tensor:
        """Copy of a workspace tap of the last forward (include/deepclr_b200.h: dclr_model_debug_copy)."""
        out = torch.empty(shape, dtype=dtype, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().dclr_model_debug_copy(self._handle, which, _lib.ptr(out),
                                                        out.numel() * out.element_size(), _lib.stream_ptr()),
                       'dclr_model_debug_copy')
        return out
