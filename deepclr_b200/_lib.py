"""ctypes loader for libdeepclr_b200.so (the C-ABI declared in include/deepclr_b200.h).

There is no fallback: if the shared library is missing or a call fails, a RuntimeError is raised.
"""
import ctypes
import os
import subprocess

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# DCLR_LIB: an alternative build of the same library (A/B kernel experiments, tools/ab.sh); default = the in-tree build
LIB_PATH = os.environ.get('DCLR_LIB') or os.path.join(_HERE, 'libdeepclr_b200.so')
_lib = None

c_int, c_float, c_void_p, c_size_t = ctypes.c_int, ctypes.c_float, ctypes.c_void_p, ctypes.c_size_t


class DclrArch(ctypes.Structure):
    """Mirror of `dclr_arch` (include/deepclr_b200.h)."""
    _fields_ = [('input_dim', c_int), ('npoint', c_int), ('num_scales', c_int), ('radii', c_float * 2),
                ('nsamples', c_int * 2), ('sa_mlp', (c_int * 3) * 2), ('knn_k', c_int), ('merge_radius', c_float),
                ('merge_mlp', c_int * 3), ('out_mlp', c_int * 5), ('out_linear', c_int * 3), ('label_type', c_int)]


_SIGNATURES = {
    'dclr_version': (ctypes.c_char_p, []),
    'dclr_last_error': (ctypes.c_char_p, []),
    'dclr_fps': (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    'dclr_fps_static': (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    'dclr_fps_static_limits': (c_int, [c_void_p, c_void_p]),
    'dclr_gather': (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    'dclr_pose_to_matrix': (c_int, [c_void_p, c_int, c_int, ctypes.c_float, c_void_p, c_void_p]),
    'dclr_pose_errors': (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    'dclr_filter_clouds': (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_float, c_float,
                                   c_void_p, c_void_p, c_void_p]),
    'dclr_subsample_clouds': (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, ctypes.c_uint64, c_int,
                                      c_void_p, c_void_p, c_void_p, c_void_p]),
    'dclr_gather_grad': (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    'dclr_group_grad': (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    'dclr_ball_query': (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_float, c_int, c_void_p, c_void_p,
                                c_int, c_void_p]),
    'dclr_group': (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    'dclr_knn': (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    'dclr_model_create': (c_int, [ctypes.POINTER(DclrArch), c_void_p, c_size_t, ctypes.POINTER(c_void_p)]),
    'dclr_model_destroy': (None, [c_void_p]),
    'dclr_model_set_numerics': (c_int, [c_void_p, c_int]),
    'dclr_model_set_fps_tie_rule': (c_int, [c_void_p, c_int]),
    'dclr_model_get_fps_tie_rule': (c_int, [c_void_p]),
    'dclr_cloud_features': (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    'dclr_forward_features': (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    'dclr_forward_pairs': (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    'dclr_workspace_bytes': (c_size_t, [c_void_p, c_int, c_int]),
    'dclr_sa_msg_fused': (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    'dclr_flow_embed_fused': (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    'dclr_pose_head_fused': (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    'dclr_model_last_launches': (c_int, [c_void_p]),
    'dclr_model_generation': (c_int, [c_void_p]),
    'dclr_model_set_timing': (c_int, [c_void_p, c_int]),
    'dclr_model_stage_count': (c_int, []),
    'dclr_model_stage_name': (ctypes.c_char_p, [c_int]),
    'dclr_model_collect_timing': (c_int, [c_void_p, c_void_p, ctypes.POINTER(c_int)]),
    'dclr_model_debug_copy': (c_int, [c_void_p, c_int, c_void_p, c_size_t, c_void_p]),
}
EXPORTS = tuple(_SIGNATURES)


def build(force: bool = False) -> str:
    """Compile the CUDA sources for sm_100a with the in-tree Makefile (nvcc cross-compiles without a GPU)."""
    src_dir = os.path.join(_HERE, 'csrc')
    if force:
        subprocess.check_call(['make', '-s', '-C', src_dir, 'clean'])
    subprocess.check_call(['make', '-s', '-j8', '-C', src_dir])
    return LIB_PATH


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f'{LIB_PATH} is missing: build it with deepclr_b200._lib.build() / '
                               f'`make -C deepclr_b200/csrc` (there is no CPU or PyTorch fallback)')
        _lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(_lib, name)
            fn.restype = res
            fn.argtypes = args
    return _lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError(f'{what} failed ({rc}): {lib().dclr_last_error().decode()}')


def ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def stream_ptr():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError('deepclr_b200: tensors must live on a CUDA device (no CPU path exists)')
