"""ctypes bindings for the CPU oracle (oracle/dclr_oracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module.
Inputs/outputs are numpy arrays in the reference's layouts (pointnet2_utils.py:12-29, 42-60, 159-177, 203-221 and
deepclr/models/deepclr.py:164-167 of the reference tree).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, 'libdclr_oracle.so')
_lib = None


def build(force: bool = False) -> str:
    """Compile the C restatement with the committed Makefile (gcc only, seconds)."""
    src = os.path.join(_HERE, 'dclr_oracle.c')
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(['make', '-s', '-C', _HERE, 'libdclr_oracle.so'])
    return _LIB_PATH


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_LIB_PATH)
        _lib.oracle_knn.restype = ctypes.c_int
        _lib.oracle_num_threads.restype = ctypes.c_int
    return _lib


def _p(a: np.ndarray):
    return a.ctypes.data_as(ctypes.c_void_p)


def _f32(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.float32)


def num_threads() -> int:
    return int(lib().oracle_num_threads())


def fps(xyz, npoint: int) -> np.ndarray:
    """Reference-faithful FPS (block-tree tie-break). xyz [B,N,3] -> idx [B,npoint] int32."""
    xyz = _f32(xyz)
    B, N, _ = xyz.shape
    out = np.zeros((B, npoint), dtype=np.int32)
    lib().oracle_fps(_p(xyz), B, N, npoint, _p(out))
    return out


def fps_lowidx(pts, npoint: int, return_ties: bool = False):
    """FPS with the lowest-index tie-break. pts [B,N,S] with S >= 3 (xyz first) -> idx [B,npoint] int32."""
    pts = _f32(pts)
    B, N, S = pts.shape
    out = np.zeros((B, npoint), dtype=np.int32)
    ties = np.zeros((B,), dtype=np.int32)
    lib().oracle_fps_lowidx(_p(pts), S, B, N, npoint, _p(out), _p(ties))
    return (out, ties) if return_ties else out


def gather(points, idx) -> np.ndarray:
    """points [B,C,N], idx [B,M] int32 -> [B,C,M]."""
    points = _f32(points)
    idx = np.ascontiguousarray(idx, dtype=np.int32)
    B, C, N = points.shape
    M = idx.shape[1]
    out = np.empty((B, C, M), dtype=np.float32)
    lib().oracle_gather(_p(points), _p(idx), B, C, N, M, _p(out))
    return out


def ball_query(radius: float, nsample: int, xyz, new_xyz, return_cnt: bool = False):
    """xyz [B,N,3], new_xyz [B,M,3] -> idx [B,M,nsample] int32 (fully padded, zeros when empty)."""
    xyz = _f32(xyz)
    new_xyz = _f32(new_xyz)
    B, N, _ = xyz.shape
    M = new_xyz.shape[1]
    idx = np.empty((B, M, nsample), dtype=np.int32)
    cnt = np.zeros((B, M), dtype=np.int32)
    lib().oracle_ball_query(_p(new_xyz), _p(xyz), B, N, M, ctypes.c_float(radius), nsample, _p(idx), _p(cnt))
    return (idx, cnt) if return_cnt else idx


def group(points, idx) -> np.ndarray:
    """points [B,C,N], idx [B,M,ns] int32 -> [B,C,M,ns]."""
    points = _f32(points)
    idx = np.ascontiguousarray(idx, dtype=np.int32)
    B, C, N = points.shape
    _, M, ns = idx.shape
    out = np.empty((B, C, M, ns), dtype=np.float32)
    lib().oracle_group(_p(points), _p(idx), B, C, N, M, ns, _p(out))
    return out


def knn(x, y, k: int, B: int) -> np.ndarray:
    """torch-cluster layout: x [B*Nx,3] candidates, y [B*Ny,3] queries -> [2, B*Ny*k] int64 (row=y id, col=x id)."""
    x = _f32(x)
    y = _f32(y)
    Nx = x.shape[0] // B
    Ny = y.shape[0] // B
    out = np.empty((2, B * Ny * k), dtype=np.int64)
    rc = lib().oracle_knn(_p(x), _p(y), B, Nx, Ny, k, _p(out[0]), _p(out[1]))
    if rc != 0:
        raise ValueError('oracle_knn: need 0 < k <= Nx')
    return out
