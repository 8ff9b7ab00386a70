"""dclr_fps_static (include/deepclr_b200.h): the farthest point sampling entry that allocates nothing -- its scratch is a
file-scope device array of the library (csrc/fps_static.cu) -- against dclr_fps, through the C-ABI.  Collected last on
purpose (the file name sorts after the other test files).

Indices are integer work: bit-exact, as everywhere in tests/test_gpu_ops.py; the gathered coordinates are copies of input
floats, so they are compared bit-exact too."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'
TIE_LOWEST, TIE_REFERENCE = 0, 1


def _clouds(B, N, seed, stride=3):
    from deepclr_b200 import synth
    pts = np.stack([synth.kitti_pair(seed + i, N)[i % 2][:, :stride] for i in range(B)]).astype(np.float32)
    return torch.from_numpy(np.ascontiguousarray(pts)).to(DEV)


def _fps(entry, pts, npoint, tie_rule):
    from deepclr_b200 import _lib
    B, N, stride = pts.shape
    idx = torch.full((B, npoint), -7, dtype=torch.int32, device=pts.device)
    new_xyz = torch.full((B, npoint, 3), float('nan'), dtype=torch.float32, device=pts.device)
    with torch.cuda.device(pts.device):
        rc = getattr(_lib.lib(), entry)(_lib.ptr(pts), stride, B, N, npoint, tie_rule, _lib.ptr(idx), _lib.ptr(new_xyz),
                                        _lib.stream_ptr())
    _lib.check(rc, entry)
    return idx, new_xyz


def _same(a, b):
    return torch.equal(a[0], b[0]) and torch.equal(a[1].view(torch.int32), b[1].view(torch.int32))


def test_limits_are_reported():
    from deepclr_b200 import _lib
    per_cloud, per_call = ctypes.c_int(0), ctypes.c_int(0)
    assert _lib.lib().dclr_fps_static_limits(ctypes.byref(per_cloud), ctypes.byref(per_call)) == 0
    assert (per_cloud.value, per_call.value) == (16384, 131072)


@pytest.mark.parametrize('tie_rule', [TIE_REFERENCE, TIE_LOWEST])
def test_two_calls_in_a_row_equal_dclr_fps(tie_rule):
    """The second call is smaller than the first, so the array still holds records of the first call behind the second
    one's: stale contents must not leak into the result.  The first call is the stated maximum (8 clouds x 16384)."""
    first, second = _clouds(8, 16384, 4100), _clouds(3, 12000, 4200, stride=4)
    want = [_fps('dclr_fps', first, 1024, tie_rule), _fps('dclr_fps', second, 1024, tie_rule)]
    got = [_fps('dclr_fps_static', first, 1024, tie_rule), _fps('dclr_fps_static', second, 1024, tie_rule)]
    torch.cuda.synchronize()
    for w, g in zip(want, got):
        assert int(g[0].min()) >= 0 and not bool(torch.isnan(g[1]).any())
        assert _same(w, g)
    # and once more in the other order, back to back on the stream, nothing in between
    again = [_fps('dclr_fps_static', second, 1024, tie_rule), _fps('dclr_fps_static', first, 1024, tie_rule)]
    assert _same(want[1], again[0]) and _same(want[0], again[1])


def test_small_clouds_take_the_resident_kernel():
    """N <= 8192 needs no scratch at all (register/shared-memory resident kernel); same entry, same results."""
    for B, N, npoint in ((5, 8192, 512), (2, 1000, 1200), (1, 1, 4)):
        pts = _clouds(B, N, 4300 + N)
        for tie_rule in (TIE_REFERENCE, TIE_LOWEST):
            assert _same(_fps('dclr_fps', pts, npoint, tie_rule), _fps('dclr_fps_static', pts, npoint, tie_rule))


def test_sizes_beyond_the_static_scratch_are_rejected():
    from deepclr_b200 import _lib
    for B, N, what in ((1, 16385, 'points per cloud'), (9, 16384, 'points per call'), (17, 8192, 'points per call')):
        pts = torch.zeros((B, N, 3), device=DEV)
        with pytest.raises(RuntimeError, match='static scratch limit'):
            _fps('dclr_fps_static', pts, 16, TIE_REFERENCE)
        assert what in _lib.lib().dclr_last_error().decode()
    with pytest.raises(RuntimeError, match='tie_rule'):
        _fps('dclr_fps_static', torch.zeros((1, 64, 3), device=DEV), 4, 5)


def test_no_device_memory_is_taken_and_the_call_can_be_captured():
    """Free device memory does not move across calls (after one call per shape has loaded the kernels), and the call is
    legal inside a CUDA graph capture, where cudaMalloc / cudaFree are not."""
    a, b = _clouds(8, 16384, 4400), _clouds(2, 9000, 4500)
    want = [_fps('dclr_fps', a, 1024, TIE_REFERENCE), _fps('dclr_fps', b, 512, TIE_REFERENCE)]
    _fps('dclr_fps_static', a, 1024, TIE_REFERENCE)
    _fps('dclr_fps_static', b, 512, TIE_REFERENCE)
    torch.cuda.synchronize()
    from deepclr_b200 import _lib
    B, N, _ = a.shape
    idx = torch.empty((B, 1024), dtype=torch.int32, device=DEV)
    new_xyz = torch.empty((B, 1024, 3), dtype=torch.float32, device=DEV)
    idx_b = torch.empty((2, 512), dtype=torch.int32, device=DEV)
    free_before = torch.cuda.mem_get_info()[0]
    for _ in range(3):
        rc = _lib.lib().dclr_fps_static(_lib.ptr(a), 3, B, N, 1024, TIE_REFERENCE, _lib.ptr(idx), _lib.ptr(new_xyz),
                                        _lib.stream_ptr())
        assert rc == 0
        rc = _lib.lib().dclr_fps_static(_lib.ptr(b), 3, 2, 9000, 512, TIE_REFERENCE, _lib.ptr(idx_b), None,
                                        _lib.stream_ptr())
        assert rc == 0
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free_before
    assert torch.equal(idx, want[0][0]) and torch.equal(idx_b, want[1][0])

    idx.fill_(-1)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        rc = _lib.lib().dclr_fps_static(_lib.ptr(a), 3, B, N, 1024, TIE_REFERENCE, _lib.ptr(idx), _lib.ptr(new_xyz),
                                        _lib.stream_ptr())
    assert rc == 0, _lib.lib().dclr_last_error().decode()
    for _ in range(2):
        idx.fill_(-1)
        graph.replay()
        torch.cuda.synchronize()
        assert torch.equal(idx, want[0][0]) and torch.equal(new_xyz, want[0][1])
