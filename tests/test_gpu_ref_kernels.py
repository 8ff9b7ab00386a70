"""Second, authoritative oracle on the GPU box: the REFERENCE's own CUDA kernels (oracle/_ref/libpointnet2_ref.so, built
from /root/reference by oracle/build_ref.sh) against this repo's kernels, bit for bit, at full KITTI size."""
import numpy as np
import pytest
import torch

from oracle import ref_cuda

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not ref_cuda.available(), reason='oracle/_ref not built')]
DEV = 'cuda:0'


@pytest.fixture(scope='module')
def p2():
    from deepclr_b200 import pointnet2
    return pointnet2


@pytest.mark.parametrize('kind,B,N,npoint', [('uniform', 3, 1000, 128), ('uniform', 2, 4096, 512),
                                             ('kitti', 4, 16384, 1024), ('modelnet', 4, 1024, 512)])
def test_fps_equals_reference_kernel(p2, kind, B, N, npoint):
    from deepclr_b200 import synth
    if kind == 'uniform':
        xyz = np.random.default_rng(N).uniform(-40, 40, (B, N, 3)).astype(np.float32)
    else:
        xyz = synth.batch(kind, B // 2, 300, n=N)[:, :, :3].copy()
    x = torch.from_numpy(xyz).to(DEV)
    want = ref_cuda.furthest_point_sample(x, npoint).cpu().numpy()
    got = p2.furthest_point_sample(x, npoint).cpu().numpy()
    # identical unless a round had an exact tie of the running maximum (the default rule is the lowest index, the
    # reference's is the tree position of sampling_gpu.cu:86-91); these seeds have none
    assert np.array_equal(got, want)


@pytest.mark.parametrize('kind,B,N,npoint', [('lattice', 2, 700, 300), ('lattice', 2, 3000, 600), ('dup', 3, 1024, 512),
                                             ('dup', 2, 5000, 700), ('dup', 2, 16384, 1024), ('dup', 1, 20000, 300),
                                             ('kitti', 16, 16384, 1024), ('kitti', 2, 40000, 2500), ('dup', 1, 70001, 600),
                                             ('lattice', 1, 100000, 300), ('kitti', 2, 131072, 1024)])
def test_fps_reference_tie_rule_is_bit_identical_to_reference_kernel(p2, reference_tie_rule, kind, B, N, npoint):
    """With dclr_set_fps_tie_rule(1) the sampled indices equal the reference kernel's on EVERY input, exact ties
    included: integer lattices and duplicated points (many bit-equal maxima), and 16 full-size KITTI-shaped scans
    (where ~1 round in 10^4 ties by chance).  All kernels: resident (N <= 2048), bucketed in shared memory
    (<= 16384), bucketed with the point state in L2 (<= 131072; one and two buckets per thread)."""
    from deepclr_b200 import synth
    rng = np.random.default_rng(N + npoint)
    if kind == 'lattice':
        xyz = rng.integers(0, 12, (B, N, 3)).astype(np.float32)
    elif kind == 'dup':
        xyz = rng.uniform(-40, 40, (B, N, 3)).astype(np.float32)
        src = rng.integers(0, N, N // 3)
        xyz[:, rng.permutation(N)[:N // 3]] = xyz[:, src]          # a third of the points are exact copies
    else:
        xyz = synth.batch('kitti', B // 2, 100, n=N)[:, :, :3].copy()
    x = torch.from_numpy(np.ascontiguousarray(xyz)).to(DEV)
    want = ref_cuda.furthest_point_sample(x, npoint)
    got = p2.furthest_point_sample(x, npoint)
    assert torch.equal(got, want), f'first mismatch at {torch.nonzero(got != want)[:3].tolist()}'


@pytest.mark.parametrize('N,M,r,ns', [(16384, 1024, 0.5, 512), (16384, 1024, 1.0, 1024), (1024, 512, 0.2, 512),
                                      (1500, 200, 0.1, 256)])
def test_ball_query_gather_group_equal_reference_kernels(p2, N, M, r, ns):
    from deepclr_b200 import synth
    if N >= 2048:
        x = synth.batch('kitti', 1, 17, n=N)
    else:
        x = synth.batch('modelnet', 1, 17, n=N)
    xyz = torch.from_numpy(x[:, :, :3].copy()).to(DEV)
    fidx = ref_cuda.furthest_point_sample(xyz, M)
    flipped = xyz.transpose(1, 2).contiguous()
    new_ref = ref_cuda.gather_operation(flipped, fidx)
    new_got = p2.gather_operation(flipped, fidx)
    assert torch.equal(new_got, new_ref)
    new_xyz = new_ref.transpose(1, 2).contiguous()
    want = ref_cuda.ball_query(r, ns, xyz, new_xyz)
    got = p2.ball_query(r, ns, xyz, new_xyz)
    assert torch.equal(got, want)
    assert torch.equal(p2.grouping_operation(flipped, got), ref_cuda.grouping_operation(flipped, want))


def test_forward_close_to_reference_cuda_composition():
    """Whole forward: this repo (fp32 mode) vs reference CUDA ops + torch fp32 dense layers (TF32 off)."""
    from oracle import model_oracle
    from tests.util import build_from_golden, golden_input, load_golden
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    g = load_golden('kitti_b1')
    model, cfg = build_from_golden(g, DEV)
    x = torch.from_numpy(golden_input(g)).to(DEV)
    y_ref = ref_cuda.forward(model.state_dict(), cfg, x).cpu().numpy()
    model.numerics = 'fp32'
    with torch.no_grad():
        y = model(x)[0].cpu().numpy()
    for b in range(y.shape[0]):
        rot, trans = model_oracle.pose_errors(y[b], y_ref[b])
        assert rot <= 0.01 and trans <= 1e-3, (rot, trans)


def test_gather_and_group_backward_equal_reference_kernels(p2):
    """autograd through gather_operation / grouping_operation (dclr_gather_grad / dclr_group_grad) against the
    reference's *_grad kernels on the same upstream gradient.  Both sides accumulate with fp32 atomicAdd, so they agree
    up to the order of the additions: integer-valued gradients make every partial sum exact -> bit-equal."""
    rng = np.random.default_rng(3)
    B, C, N, M, ns = 2, 5, 700, 90, 24
    feats = torch.from_numpy(rng.normal(size=(B, C, N)).astype(np.float32)).to(DEV).requires_grad_(True)
    gidx = torch.from_numpy(rng.integers(0, N, (B, M)).astype(np.int32)).to(DEV)
    qidx = torch.from_numpy(rng.integers(0, 60, (B, M, ns)).astype(np.int32)).to(DEV)     # many repeats per slot
    up1 = torch.from_numpy(rng.integers(-8, 9, (B, C, M)).astype(np.float32)).to(DEV)
    up2 = torch.from_numpy(rng.integers(-8, 9, (B, C, M, ns)).astype(np.float32)).to(DEV)
    g1, = torch.autograd.grad(p2.gather_operation(feats, gidx), feats, up1)
    g2, = torch.autograd.grad(p2.grouping_operation(feats, qidx), feats, up2)
    assert torch.equal(g1, ref_cuda.gather_operation_grad(up1, gidx, N))
    assert torch.equal(g2, ref_cuda.grouping_operation_grad(up2, qidx, N))
    # real-valued gradients: equal to a float64 scatter-add within fp32 rounding
    up3 = torch.from_numpy(rng.normal(size=(B, C, M, ns)).astype(np.float32)).to(DEV)
    g3, = torch.autograd.grad(p2.grouping_operation(feats, qidx), feats, up3)
    want = torch.zeros((B, C, N), dtype=torch.float64, device=DEV)
    want.scatter_add_(2, qidx.long().view(B, 1, M * ns).expand(-1, C, -1), up3.double().view(B, C, M * ns))
    torch.testing.assert_close(g3.double(), want, rtol=1e-5, atol=1e-5)
