"""GPU parity of the C-ABI point ops (through the pointnet2 / torch_cluster drop-ins) against the CPU oracle:
indices bit-exact.  Run on the B200 box:  python -m pytest tests -m gpu"""
import numpy as np
import pytest
import torch

from oracle import ops
from tests.util import load_golden

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


@pytest.fixture(scope='module')
def p2():
    from deepclr_b200 import pointnet2
    return pointnet2


def test_golden_ops(p2):
    from deepclr_b200.torch_cluster import knn
    g = load_golden('ops_small')
    xyz, feats = _t(g['xyz']), _t(g['feats'])
    fps = p2.furthest_point_sample(xyz, 64)
    assert fps.dtype == torch.int32 and np.array_equal(fps.cpu().numpy(), g['fps_low'])
    assert np.array_equal(p2.furthest_point_sample(_t(g['small']), 128).cpu().numpy(), g['small_fps_low'])
    new_xyz = _t(g['new_xyz'])
    idx = p2.ball_query(0.35, 16, xyz, new_xyz)
    assert np.array_equal(idx.cpu().numpy(), g['ball_idx'])
    assert np.array_equal(p2.gather_operation(feats, fps).cpu().numpy(), g['gathered'])
    assert np.array_equal(p2.grouping_operation(feats, idx).cpu().numpy(), g['grouped'])
    out = knn(xyz[:2].reshape(-1, 3), new_xyz[:2].reshape(-1, 3), 9,
              torch.arange(2, device=DEV).repeat_interleave(300), torch.arange(2, device=DEV).repeat_interleave(64))
    assert out.dtype == torch.int64 and np.array_equal(out.cpu().numpy(), g['knn'])


@pytest.mark.parametrize('B,N,npoint', [(1, 1, 1), (2, 33, 33), (3, 96, 1024), (2, 1000, 100), (2, 1024, 512),
                                        (1, 2500, 300), (2, 5000, 257), (1, 9000, 128), (2, 16384, 1024),
                                        (1, 20000, 64), (2, 33333, 150), (1, 70001, 200), (1, 140000, 40)])
def test_fps_matches_oracle(p2, B, N, npoint):
    rng = np.random.default_rng(N + npoint)
    xyz = rng.uniform(-40, 40, (B, N, 3)).astype(np.float32)
    want, ties = ops.fps_lowidx(xyz, npoint, return_ties=True)
    got = p2.furthest_point_sample(_t(xyz), npoint).cpu().numpy()
    assert np.array_equal(got, want), f'first mismatch at {np.argwhere(got != want)[:3]} (tie rounds per cloud: {ties})'
    if npoint <= N:
        assert np.array_equal(got, ops.fps(xyz, npoint)) or ties.sum() > 0   # reference tie rule agrees w/o ties


def test_fps_exact_ties_take_lowest_index(p2):
    # integer lattice: many exactly equal distances; the stated rule is the globally lowest index
    g = np.stack(np.meshgrid(np.arange(8), np.arange(8), np.arange(4), indexing='ij'), -1).reshape(1, -1, 3)
    xyz = g.astype(np.float32)
    want, ties = ops.fps_lowidx(xyz, 100, return_ties=True)
    assert ties[0] > 0
    assert np.array_equal(p2.furthest_point_sample(_t(xyz), 100).cpu().numpy(), want)


@pytest.mark.parametrize('N,npoint', [(1, 1), (7, 7), (100, 60), (700, 300), (1024, 400), (1500, 500), (2048, 64),
                                      (3000, 500), (6000, 300), (16384, 256), (17000, 64)])
def test_fps_reference_tie_rule_matches_faithful_oracle(p2, N, npoint):
    """dclr_set_fps_tie_rule(1): the block-tree order of sampling_gpu.cu:113-203 (restated in oracle_fps, block size
    opt_n_threads(N) incl. the non-power-of-two N < 1024 cases) on integer lattices where most rounds tie."""
    import deepclr_b200
    rng = np.random.default_rng(N)
    xyz = rng.integers(0, 10, (2, N, 3)).astype(np.float32)
    _, ties = ops.fps_lowidx(xyz, npoint, return_ties=True)
    assert N < 100 or ties.sum() > 0
    deepclr_b200.set_fps_tie_rule('reference')
    try:
        assert deepclr_b200.get_fps_tie_rule() == 'reference'
        got = p2.furthest_point_sample(_t(xyz), npoint).cpu().numpy()
    finally:
        deepclr_b200.set_fps_tie_rule('lowest_index')
    assert np.array_equal(got, ops.fps(xyz, npoint))


def test_fps_point_stride_4_and_new_xyz():
    from deepclr_b200 import _lib
    rng = np.random.default_rng(5)
    pts = rng.normal(size=(2, 3000, 4)).astype(np.float32)
    x = _t(pts)
    idx = torch.empty((2, 200), dtype=torch.int32, device=DEV)
    nx = torch.empty((2, 200, 3), dtype=torch.float32, device=DEV)
    _lib.check(_lib.lib().dclr_fps(_lib.ptr(x), 4, 2, 3000, 200, 0, _lib.ptr(idx), _lib.ptr(nx), _lib.stream_ptr()), 'fps')
    want = ops.fps_lowidx(pts, 200)
    assert np.array_equal(idx.cpu().numpy(), want)
    assert np.array_equal(nx.cpu().numpy(), np.stack([pts[b, want[b], :3] for b in range(2)]))


@pytest.mark.parametrize('B,N,M,r,ns', [(1, 10, 4, 0.5, 32), (2, 777, 100, 0.2, 16), (2, 4096, 256, 0.1, 64),
                                        (1, 16384, 1024, 1.0, 128), (2, 5000, 300, 3.0, 8), (1, 2048, 70, 100.0, 4096)])
def test_ball_query_matches_oracle(p2, B, N, M, r, ns):
    rng = np.random.default_rng(N + M)
    xyz = rng.uniform(-1, 1, (B, N, 3)).astype(np.float32)
    new_xyz = xyz[:, rng.permutation(N)[:M]].copy()
    new_xyz[0, 0] = 77.0                                     # empty ball (unless r is huge)
    want = ops.ball_query(r, ns, xyz, new_xyz)
    got = p2.ball_query(r, ns, _t(xyz), _t(new_xyz)).cpu().numpy()
    assert np.array_equal(got, want)


@pytest.mark.parametrize('N,M,r,ns', [(16384, 1024, 0.5, 512), (16384, 1024, 1.0, 1024), (40000, 500, 2.5, 64),
                                      (131072, 256, 0.7, 32)])
def test_ball_query_grid_path_lidar_and_clustered(p2, N, M, r, ns):
    """Large clouds go through the uniform-grid kernels (ballgrid.cu): KITTI-shaped scans (dense rings near the sensor,
    sparse far away), a degenerate cloud with most points in one cell, and centres outside the bounding box."""
    from deepclr_b200 import synth
    rng = np.random.default_rng(N + ns)
    lidar = synth._lidar_scan(np.random.Generator(np.random.PCG64(7)), N)[0].astype(np.float32)
    clustered = rng.normal(0, 0.3, (N, 3)).astype(np.float32)
    clustered[: N // 8] = rng.uniform(-50, 50, (N // 8, 3)).astype(np.float32)
    xyz = np.stack([lidar, clustered])
    new_xyz = xyz[:, rng.permutation(N)[:M]].copy()
    new_xyz[0, 0] = (500.0, -500.0, 3.0)                     # outside the box: empty
    new_xyz[1, 1] = xyz[1].max(0) + 0.5 * r                  # just outside a corner
    want = ops.ball_query(r, ns, xyz, new_xyz)
    got = p2.ball_query(r, ns, _t(xyz), _t(new_xyz)).cpu().numpy()
    assert np.array_equal(got, want)


def test_ball_query_stride4_two_radii_compact():
    from deepclr_b200 import _lib
    rng = np.random.default_rng(9)
    pts = rng.uniform(-2, 2, (2, 3000, 4)).astype(np.float32)
    ctr = pts[:, :64, :3].copy()
    x, c = _t(pts), _t(ctr)
    for r, ns in ((0.3, 16), (0.6, 32)):
        idx = torch.full((2, 64, ns), -7, dtype=torch.int32, device=DEV)
        cnt = torch.empty((2, 64), dtype=torch.int32, device=DEV)
        _lib.check(_lib.lib().dclr_ball_query(_lib.ptr(c), _lib.ptr(x), 4, 2, 3000, 64, r, ns, _lib.ptr(idx),
                                              _lib.ptr(cnt), 0, _lib.stream_ptr()), 'bq')
        w_idx, w_cnt = ops.ball_query(r, ns, pts[:, :, :3], ctr, return_cnt=True)
        idx, cnt = idx.cpu().numpy(), cnt.cpu().numpy()
        assert np.array_equal(cnt, w_cnt)
        for b in range(2):
            for i in range(64):
                assert np.array_equal(idx[b, i, :cnt[b, i]], w_idx[b, i, :cnt[b, i]])
                assert (idx[b, i, cnt[b, i]:] == -7).all()          # pad=0 leaves the tail untouched


@pytest.mark.parametrize('B,Nx,Ny,k', [(1, 20, 5, 20), (2, 512, 512, 30), (3, 1024, 1024, 20), (1, 3000, 100, 100),
                                       (2, 1025, 64, 7), (1, 16384, 300, 20), (2, 5000, 33, 128)])
def test_knn_matches_oracle(B, Nx, Ny, k):
    from deepclr_b200.torch_cluster import knn
    rng = np.random.default_rng(Nx + k)
    x = rng.normal(size=(B * Nx, 3)).astype(np.float32)
    y = rng.normal(size=(B * Ny, 3)).astype(np.float32)
    x[1] = x[0]                                              # exact tie -> lower index first
    bx = torch.arange(B, device=DEV).repeat_interleave(Nx)
    by = torch.arange(B, device=DEV).repeat_interleave(Ny)
    got = knn(_t(x), _t(y), k, bx, by).cpu().numpy()
    assert np.array_equal(got, ops.knn(x, y, k, B))
    with pytest.raises(RuntimeError):
        knn(_t(x[:4]), _t(y[:4]), 5)


@pytest.mark.parametrize('order', ['towards', 'away', 'ties'])
def test_knn_general_kernel_on_ordered_candidates(order):
    """The general kernel prunes a tile by the current k-th distance: candidates sorted TOWARDS the queries make every
    tile overflow its merge buffer (fallback path), sorted AWAY nothing after the first tile survives, and a cloud of
    few distinct positions is all ties (lower index first, as torch-cluster's insertion does)."""
    from deepclr_b200.torch_cluster import knn
    rng = np.random.default_rng(3)
    Nx, Ny, k = 6000, 40, 24
    y = (rng.normal(size=(Ny, 3)) * 0.01).astype(np.float32)
    x = rng.normal(size=(Nx, 3)).astype(np.float32)
    r = np.linalg.norm(x, axis=1)
    if order == 'towards':
        x = x[np.argsort(-r)]
    elif order == 'away':
        x = x[np.argsort(r)]
    else:
        x = x[rng.integers(0, 7, Nx)]
    got = knn(_t(x), _t(y), k).cpu().numpy()
    assert np.array_equal(got, ops.knn(x, y, k, 1))


def test_gather_group_shapes_and_backward(p2):
    rng = np.random.default_rng(11)
    feats = torch.tensor(rng.normal(size=(2, 6, 50)).astype(np.float32), device=DEV, requires_grad=True)
    idx = torch.tensor(rng.integers(0, 50, (2, 9, 4)).astype(np.int32), device=DEV)
    out = p2.grouping_operation(feats, idx)
    assert np.array_equal(out.detach().cpu().numpy(), ops.group(feats.detach().cpu().numpy(), idx.cpu().numpy()))
    out.sum().backward()
    cnt = np.zeros((2, 50))
    for b in range(2):
        np.add.at(cnt[b], idx[b].cpu().numpy().ravel(), 1)
    np.testing.assert_allclose(feats.grad.cpu().numpy(), np.repeat(cnt[:, None, :], 6, 1))


def test_full_size_properties(p2):
    """BASELINE sizes (16384-point KITTI-shaped clouds): size-independent properties instead of the slow oracle."""
    from deepclr_b200 import synth
    x = synth.batch('kitti', 2, 100)                                     # [4,16384,4]
    xyz = _t(x[:, :, :3])
    idx = p2.furthest_point_sample(xyz, 1024).cpu().numpy()
    assert (idx[:, 0] == 0).all()
    for b in range(4):
        assert len(set(idx[b].tolist())) == 1024                        # distinct points -> distinct samples
        # greedy property: the min-distance of each new sample to the previous ones is non-increasing
        p = x[b, idx[b], :3].astype(np.float64)
        d = np.full(1024, np.inf)
        mins = []
        for j in range(1, 1024):
            d = np.minimum(d, ((p - p[j - 1]) ** 2).sum(1))
            mins.append(d[j])
        assert all(mins[i] >= mins[i + 1] - 1e-3 for i in range(len(mins) - 1))
    new_xyz = torch.stack([xyz[b, torch.from_numpy(idx[b]).long().to(DEV)] for b in range(4)])
    bq = p2.ball_query(1.0, 1024, xyz, new_xyz).cpu().numpy()
    pts = x[:, :, :3]
    for b in range(4):
        d2 = ((pts[b, bq[b, :64]] - pts[b, idx[b, :64]][:, None, :]) ** 2).sum(-1)
        assert (d2 < 1.0 + 1e-4).all()                                   # every returned index is inside the ball
        assert (bq[b, :, 0] <= idx[b]).all()                             # the centre itself is a hit; first hit <= it
        row = bq[b, 5]
        uniq = row[:len(set(row.tolist()))]
        assert (np.diff(uniq) > 0).all()                                 # ascending index order


@pytest.mark.parametrize('lt_name', ['POSE3D_EULER', 'POSE3D_QUAT', 'POSE3D_DUAL_QUAT'])
def test_device_to_matrix_and_kitti_errors_match_host(lt_name):
    """Batched on-device label -> matrix and KITTI errors (dclr_pose_to_matrix / dclr_pose_errors, float64) against the
    host restatement of LabelType.to_matrix (deepclr/data/labels.py:80-101) and of the error definitions
    (deepclr/evaluation/metrics.py:8-49).  Tolerance 1e-12: same float64 formulas."""
    from deepclr_b200.data import device as dev_post
    from deepclr_b200.data.labels import LabelType
    lt = getattr(LabelType, lt_name)
    rng = np.random.default_rng(11)
    B = 200
    mats = []
    for _ in range(2 * B):
        ang = rng.uniform(-0.3, 0.3, 3)
        m = np.eye(4)
        cx, sx, cy, sy, cz, sz = np.cos(ang[0]), np.sin(ang[0]), np.cos(ang[1]), np.sin(ang[1]), np.cos(ang[2]), np.sin(ang[2])
        m[:3, :3] = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]]) @ np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]]) \
            @ np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
        m[:3, 3] = rng.uniform(-2, 2, 3)
        mats.append(m)
    labels = np.stack([lt.from_matrix(m) for m in mats]).astype(np.float32)
    labels[:, :] *= rng.uniform(0.5, 2.0, (2 * B, 1)).astype(np.float32) if lt_name == 'POSE3D_DUAL_QUAT' else 1.0
    want = np.stack([lt.to_matrix(l) for l in labels])
    got = dev_post.to_matrix(_t(labels), lt)
    np.testing.assert_allclose(got.cpu().numpy(), want, rtol=0, atol=1e-12)
    got_s = dev_post.to_matrix(_t(labels * 3.0), lt, scale=3.0)
    np.testing.assert_allclose(got_s.cpu().numpy(), np.stack([lt.to_matrix(l * 3.0, 3.0) for l in labels]), atol=1e-6)
    rot, trans = dev_post.kitti_errors(got[:B], got[B:])

    def host_err(a, b):
        out = []
        for p, q in ((a, b), (b, a)):
            d = p @ np.linalg.inv(q)
            out.append((np.arccos(max(min(0.5 * (d[0, 0] + d[1, 1] + d[2, 2] - 1.0), 1.0), -1.0)),
                        np.linalg.norm(d[:3, 3])))
        return min(out[0][0], out[1][0]), min(out[0][1], out[1][1])
    ref = np.array([host_err(want[i], want[B + i]) for i in range(B)])
    # acos amplifies rounding near 0 / pi; these rotations are well away from both
    np.testing.assert_allclose(rot.cpu().numpy(), ref[:, 0], atol=1e-9)
    np.testing.assert_allclose(trans.cpu().numpy(), ref[:, 1], atol=1e-12)


@pytest.mark.parametrize('shape,mp', [(8, 0), (16, 0), (8, 1), (8, 2), (8, 3), (8, 4), (8, 6), (16, 1), (16, 2), (16, 4), (32, 0),
                                      (-2, 2), (-2, 3), (-2, 4), (-2, 6)])
def test_fps_multipick_variants_are_exact(p2, shape, mp, monkeypatch):
    """Every warps-per-cloud x samples-per-round variant of the bucketed kernels (fps_tmem.cu: minima in tensor memory,
    mp >= 1 samples per round; fps_bucket.cu: mp = 0, registers; DCLR_FPS_SHAPE / DCLR_FPS_MP are read per call) returns the indices of the serial algorithm (sampling_gpu.cu:113-208), under both tie
    rules: KITTI-shaped scans, a cloud with a third of its points duplicated and an integer lattice (most rounds tie), a
    ragged N, and npoint > N (all minima reach 0: the strict-> start of the reference, sampling_gpu.cu:119-137)."""
    import deepclr_b200
    from deepclr_b200 import synth
    monkeypatch.setenv('DCLR_FPS_SHAPE', str(abs(shape) if shape > 0 else 8))
    monkeypatch.setenv('DCLR_FPS_MP', str(mp))
    monkeypatch.setenv('DCLR_FPS_CLUSTER', '1' if shape < 0 else '0')       # shape -2: a cluster of two CTAs per cloud (fps_cluster.cu)
    rng = np.random.default_rng(7)
    dup = rng.uniform(-40, 40, (2, 16384, 3)).astype(np.float32)
    dup[:, rng.permutation(16384)[:5000]] = dup[:, rng.integers(0, 16384, 5000)]
    cases = [(np.stack([synth.kitti_pair(300 + i, 16384)[i % 2][:, :3] for i in range(4)]), 1024),
             (dup, 700),
             (rng.integers(0, 14, (2, 12000, 3)).astype(np.float32), 400),
             (rng.uniform(-30, 30, (3, 9001, 3)).astype(np.float32), 333),
             (rng.uniform(-1, 1, (1, 8200, 3)).astype(np.float32), 8300)]
    monkeypatch.setenv('DCLR_FPS_BUCKET_MIN', '2048')
    cases.append((rng.uniform(-5, 5, (2, 3000, 3)).astype(np.float32), 256))      # the 4096- and 8192-record classes
    cases.append((rng.uniform(-5, 5, (2, 6000, 3)).astype(np.float32), 256))
    for xyz, npoint in cases:
        want_low = ops.fps_lowidx(xyz, npoint)
        got = p2.furthest_point_sample(_t(xyz), npoint).cpu().numpy()
        assert np.array_equal(got, want_low), (xyz.shape, npoint, int((got != want_low).sum()))
        deepclr_b200.set_fps_tie_rule('reference')
        got = p2.furthest_point_sample(_t(xyz), npoint).cpu().numpy()
        deepclr_b200.set_fps_tie_rule('lowest_index')
        assert np.array_equal(got, ops.fps(xyz, npoint)), (xyz.shape, npoint)
