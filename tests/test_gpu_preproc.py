"""GPU parity of the on-device pre-processing (dclr_filter_clouds / dclr_subsample_clouds, through the C ABI) against
the fixtures made by the reference's own transform classes (tests/golden/preproc_small.npz) and the numpy restatement
in oracle/preproc.py: bit-exact rows, counts and chosen indices."""
import numpy as np
import pytest
import torch

from oracle import preproc
from tests.util import load_golden

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


@pytest.fixture(scope='module')
def dev_ops():
    from deepclr_b200.data import device
    return device


def _rows(rc, b):
    o, c = int(rc.offsets[b]), int(rc.counts[b])
    return rc.points[o:o + c].cpu().numpy()


def test_filter_matches_reference_transform_goldens(dev_ops):
    g = load_golden('preproc_small')
    clouds = [g[f'cloud_{b}'] for b in range(int(g['num_clouds']))]
    for ci in range(int(g['num_cfgs'])):
        nth, start, rmin, rmax, dim = g[f'cfg_{ci}']
        rc = dev_ops.RaggedClouds.from_list(clouds, DEV)
        out = dev_ops.filter_clouds(rc, int(nth), int(start), float(rmin), float(rmax), out_dim=int(dim))
        for b in range(len(clouds)):
            want = g[f'out_{ci}_{b}']
            got = _rows(out, b)
            assert got.shape == want.shape, (ci, b, got.shape, want.shape)
            assert np.array_equal(got, want, equal_nan=True), (ci, b)


@pytest.mark.parametrize('C,nth,start,rmin,rmax', [(4, 1, 0, 0.0, np.inf), (4, 2, 0, 2.0, 50.0), (4, 2, 1, 0.0, 30.0),
                                                   (3, 1, 0, 1.0, 60.0), (6, 4, 3, 2.0, np.inf)])
def test_filter_full_size_scans_bit_exact(dev_ops, C, nth, start, rmin, rmax):
    rng = np.random.Generator(np.random.PCG64(11))
    sizes = [123457, 60000, 16384, 2048, 2049, 0, 1, 98765]
    clouds = [np.concatenate([rng.uniform(-80, 80, (n, 2)), rng.uniform(-3, 3, (n, C - 2))], 1).astype(np.float32)
              for n in sizes]
    rc = dev_ops.RaggedClouds.from_list(clouds, DEV)
    out_dim = C if C != 6 else 3
    out = dev_ops.filter_clouds(rc, nth, start, rmin, rmax, out_dim=out_dim)
    counts = out.counts.cpu().numpy()
    for b, c in enumerate(clouds):
        want = preproc.filter_cloud(c, nth, start, rmin, rmax, out_dim)
        assert counts[b] == want.shape[0], (b, counts[b], want.shape)
        assert np.array_equal(_rows(out, b), want)


def test_filter_cloud_longer_than_max_points_is_reported(dev_ops):
    """max_points sizes the grid and the tile status array: a longer cloud is not processed, its count reads -1, the
    other clouds of the call are unaffected."""
    from deepclr_b200 import _lib
    rng = np.random.Generator(np.random.PCG64(3))
    c = rng.uniform(-80, 80, (50000 + 3000, 4)).astype(np.float32)
    pts = torch.from_numpy(c).to(DEV)
    off = torch.tensor([0, 50000, 53000], dtype=torch.int32, device=DEV)
    out = torch.empty_like(pts)
    cnt = torch.empty(2, dtype=torch.int32, device=DEV)
    _lib.check(_lib.lib().dclr_filter_clouds(_lib.ptr(pts), _lib.ptr(off), 2, 3000, 4, 4, 2, 0, 2.0, 50.0,
                                             _lib.ptr(out), _lib.ptr(cnt), _lib.stream_ptr()), 'dclr_filter_clouds')
    want = preproc.filter_cloud(c[50000:], 2, 0, 2.0, 50.0)
    assert int(cnt[0]) == -1
    assert int(cnt[1]) == want.shape[0] and np.array_equal(out[50000:50000 + int(cnt[1])].cpu().numpy(), want)


def test_filter_argument_errors(dev_ops):
    from deepclr_b200 import _lib
    z = torch.zeros((8, 4), device=DEV)
    off = torch.tensor([0, 8], dtype=torch.int32, device=DEV)
    cnt = torch.empty(1, dtype=torch.int32, device=DEV)
    L = _lib.lib()
    assert L.dclr_filter_clouds(_lib.ptr(z), _lib.ptr(off), 1, 8, 4, 4, 0, 0, 0.0, 1.0, _lib.ptr(z), _lib.ptr(cnt), None) < 0
    assert L.dclr_filter_clouds(_lib.ptr(z), _lib.ptr(off), 1, 8, 4, 5, 1, 0, 0.0, 1.0, _lib.ptr(z), _lib.ptr(cnt), None) < 0
    assert L.dclr_filter_clouds(_lib.ptr(z), _lib.ptr(off), 1, 8, 4, 4, 2, 2, 0.0, 1.0, _lib.ptr(z), _lib.ptr(cnt), None) < 0
    assert b'nth' in L.dclr_last_error()
    assert L.dclr_subsample_clouds(_lib.ptr(z), _lib.ptr(off), None, 1, 4, 5, 4, 0, 1, _lib.ptr(z), None, None, None) < 0


@pytest.mark.parametrize('C', [4, 3])
def test_subsample_rows_match_restated_permutation(dev_ops, C):
    rng = np.random.Generator(np.random.PCG64(5))
    sizes = [700, 256, 0, 1, 300, 257, 4096]
    n_out = 256
    clouds = [rng.normal(size=(n, C)).astype(np.float32) for n in sizes]
    rc = dev_ops.RaggedClouds.from_list(clouds, DEV)
    x, perm, short = dev_ops.subsample_clouds(rc, n_out, seed=0x1234567890abcdef, return_perm=True, return_short=True)
    x, perm = x.cpu().numpy(), perm.cpu().numpy()
    assert int(short) == 2                                     # the empty cloud and the 1-point cloud
    for b, c in enumerate(clouds):
        want = preproc.subsample_rows(sizes[b], n_out, 0x1234567890abcdef, b)
        assert np.array_equal(perm[b], want), b
        if sizes[b] == 0:
            assert not x[b].any()
            continue
        assert np.array_equal(x[b], c[perm[b]])
        if sizes[b] >= n_out:
            assert len(np.unique(perm[b])) == n_out            # distinct rows
    assert np.array_equal(perm[1], np.arange(n_out))           # equal size: order kept (base.py:130-135)
    x2 = dev_ops.subsample_clouds(rc, n_out, seed=0x1234567890abcdef, ordered_if_equal=False).cpu().numpy()
    assert sorted(map(tuple, x2[1])) == sorted(map(tuple, clouds[1])) and not np.array_equal(x2[1], clouds[1])


def test_subsample_full_size_properties(dev_ops):
    """KITTI-sized: 32 scans of ~60 k points -> [32, 16384, 4]; distinct rows of the right cloud, roughly uniform."""
    rng = np.random.Generator(np.random.PCG64(9))
    sizes = rng.integers(55000, 65000, 32).tolist()
    clouds = [rng.uniform(-80, 80, (n, 4)).astype(np.float32) for n in sizes]
    rc = dev_ops.RaggedClouds.from_list(clouds, DEV)
    x, perm = dev_ops.subsample_clouds(rc, 16384, seed=77, return_perm=True)
    x, perm = x.cpu().numpy(), perm.cpu().numpy()
    x_other = dev_ops.subsample_clouds(rc, 16384, seed=78).cpu().numpy()
    assert not np.array_equal(x, x_other)
    for b, c in enumerate(clouds):
        assert perm[b].min() >= 0 and perm[b].max() < sizes[b]
        assert len(np.unique(perm[b])) == 16384
        assert np.array_equal(x[b], c[perm[b]])
        # uniform choice: mean index n/2 +- 5 sigma (sigma of the mean of 16384 draws ~ n / sqrt(12 * 16384))
        assert abs(perm[b].mean() - sizes[b] / 2) < 5 * sizes[b] / np.sqrt(12 * 16384)
        # every tenth of the cloud gets its share (binomial, 6 sigma)
        hist = np.histogram(perm[b], bins=10, range=(0, sizes[b]))[0]
        assert np.all(np.abs(hist - 1638.4) < 6 * np.sqrt(1638.4 * 0.9))


def test_prepare_pairs_feeds_the_forward(dev_ops):
    """raw scans -> filter -> subsample -> x [2B, N, C], templates first; same poses as the forward on the host-built
    batch with the same rows."""
    from tests.util import build_from_golden
    g = load_golden('modelnet_b2')
    model, cfg = build_from_golden(g, DEV)
    rng = np.random.Generator(np.random.PCG64(21))
    N = 1024
    tmpl = [rng.uniform(-1, 1, (n, 3)).astype(np.float32) for n in (1500, 1024)]
    src = [rng.uniform(-1, 1, (n, 3)).astype(np.float32) for n in (1100, 2000)]
    x = dev_ops.prepare_pairs(tmpl, src, N, DEV, input_dim=3, seed=5)
    assert x.shape == (4, N, 3)
    rows = [preproc.subsample_rows(c.shape[0], N, 5, b) for b, c in enumerate(tmpl + src)]
    x_host = np.stack([c[r] for c, r in zip(tmpl + src, rows)])
    assert np.array_equal(x.cpu().numpy(), x_host)
    with torch.no_grad():
        y, _, _ = model(x)
        y2, _, _ = model(torch.from_numpy(x_host).to(DEV))
    assert torch.equal(y, y2)


def test_stack_on_device_equalises_sizes():
    from deepclr_b200.models.base import ModelInferenceHelper
    rng = np.random.Generator(np.random.PCG64(2))
    t = torch.from_numpy(rng.normal(size=(900, 4)).astype(np.float32)).to(DEV)
    s = torch.from_numpy(rng.normal(size=(1300, 4)).astype(np.float32)).to(DEV)
    x = ModelInferenceHelper.stack(t, s)
    assert x.shape == (2, 900, 4) and torch.equal(x[0], t)
    got = {tuple(r) for r in x[1].cpu().numpy().tolist()}
    assert len(got) == 900 and got <= {tuple(r) for r in s.cpu().numpy().tolist()}


def test_raw_scan_pipeline_equals_prepare_then_forward(dev_ops):
    """RawScanPipeline (H2D of raw ragged scans -> filter -> subsample -> forward, double-buffered) gives the poses of
    the same steps done one by one."""
    from deepclr_b200.models import RawScanPipeline
    from tests.util import build_from_golden
    g = load_golden('kitti_small_b2')
    model, cfg = build_from_golden(g, DEV)
    N = int(g['n'])
    rng = np.random.Generator(np.random.PCG64(31))
    batches = []
    for _ in range(5):
        lengths = rng.integers(3 * N, 4 * N, 4).tolist()
        scans = [np.concatenate([rng.uniform(-40, 40, (n, 2)), rng.uniform(-2, 1, (n, 1)), rng.uniform(0, 1, (n, 1))],
                                1).astype(np.float32) for n in lengths]
        batches.append((torch.from_numpy(np.concatenate(scans)).pin_memory(), lengths, scans))
    pipe = RawScanPipeline(model, N, nth=2, start=0, min_range=1.0, max_range=39.0, seed=100)
    ys = pipe.run_all((b[0], b[1]) for b in batches)
    assert len(ys) == 5
    for i, (stage, lengths, scans) in enumerate(batches):
        rc = dev_ops.filter_clouds(dev_ops.RaggedClouds.from_list(scans, DEV), 2, 0, 1.0, 39.0, out_dim=4)
        assert int(rc.counts.min()) >= N
        x = dev_ops.subsample_clouds(rc, N, seed=100 + i)
        with torch.no_grad():
            y = model(x)[0]
        assert torch.equal(ys[i], y.cpu()), i
    assert pipe.h2d_bytes == sum(b[0].numel() * 4 + 5 * 4 for b in batches)


def test_filter_random_ragged_batches(dev_ops):
    """40 random ragged batches (empty clouds, sizes around the 2048-point tile edge, C = 3..6, random nth / start /
    range, +-inf and NaN coordinates): rows and counts bit-equal to the restated reference transforms."""
    rng = np.random.Generator(np.random.PCG64(2024))
    for trial in range(40):
        B = int(rng.integers(1, 24))
        C = int(rng.integers(3, 7))
        out_dim = int(rng.integers(3, C + 1))
        sizes = [int(rng.choice([0, 1, 2047, 2048, 2049, 4096, int(rng.integers(0, 9000))])) for _ in range(B)]
        clouds = []
        for n in sizes:
            c = rng.uniform(-70, 70, (n, C)).astype(np.float32)
            if n > 10:
                c[rng.integers(0, n, 3), rng.integers(0, 2, 3)] = [np.nan, np.inf, -np.inf]
            clouds.append(c)
        nth = int(rng.integers(1, 5))
        start = int(rng.integers(0, nth))
        rmin = float(rng.choice([0.0, 1.5, 10.0]))
        rmax = float(rng.choice([np.inf, 45.0, 69.0]))
        out = dev_ops.filter_clouds(dev_ops.RaggedClouds.from_list(clouds, DEV), nth, start, rmin, rmax, out_dim=out_dim)
        counts = out.counts.cpu().numpy()
        for b, c in enumerate(clouds):
            want = preproc.filter_cloud(c, nth, start, rmin, rmax, out_dim)
            assert counts[b] == want.shape[0], (trial, b)
            assert np.array_equal(_rows(out, b), want, equal_nan=True), (trial, b)


def test_kernels_write_only_what_they_own():
    """Own bounds check (compute-sanitizer is not available on the pool): outputs pre-filled with a sentinel; the
    filter may touch only the first counts[b] rows of every cloud's slot, the subsample only its [B, N, C] block; guard
    rows behind both stay intact."""
    from deepclr_b200 import _lib
    rng = np.random.Generator(np.random.PCG64(77))
    S = -12345.0
    for C, out_dim in ((4, 4), (5, 3)):
        sizes = [5000, 0, 2048, 1, 4097, 333, 2049]
        clouds = [rng.uniform(-60, 60, (n, C)).astype(np.float32) for n in sizes]
        off_h = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
        total, guard = int(off_h[-1]), 4096
        pts = torch.from_numpy(np.concatenate(clouds)).to(DEV)
        off = torch.from_numpy(off_h).to(DEV)
        out = torch.full((total + guard, out_dim), S, device=DEV)
        cnt = torch.full((len(sizes) + 8,), -7, dtype=torch.int32, device=DEV)
        _lib.check(_lib.lib().dclr_filter_clouds(_lib.ptr(pts), _lib.ptr(off), len(sizes), max(sizes), C, out_dim, 3, 1,
                                                 2.0, 50.0, _lib.ptr(out), _lib.ptr(cnt), _lib.stream_ptr()), 'filter')
        o, c = out.cpu().numpy(), cnt.cpu().numpy()
        assert (c[len(sizes):] == -7).all() and (o[total:] == S).all()
        for b, cl in enumerate(clouds):
            want = preproc.filter_cloud(cl, 3, 1, 2.0, 50.0, out_dim)
            assert c[b] == want.shape[0]
            assert np.array_equal(o[off_h[b]:off_h[b] + c[b]], want)
            assert (o[off_h[b] + c[b]:off_h[b + 1]] == S).all(), (C, b)
        n_out = 300
        x = torch.full((len(sizes) * n_out + guard, out_dim), S, device=DEV)
        perm = torch.full((len(sizes) * n_out + guard,), -9, dtype=torch.int32, device=DEV)
        _lib.check(_lib.lib().dclr_subsample_clouds(_lib.ptr(out), _lib.ptr(off), _lib.ptr(cnt), len(sizes), out_dim,
                                                    out_dim, n_out, 5, 1, _lib.ptr(x), _lib.ptr(perm), None,
                                                    _lib.stream_ptr()), 'subsample')
        assert (x[len(sizes) * n_out:] == S).all() and (perm[len(sizes) * n_out:] == -9).all()
        assert not (x[:len(sizes) * n_out] == S).any()
        p = perm[:len(sizes) * n_out].view(len(sizes), n_out).cpu().numpy()
        for b in range(len(sizes)):
            assert p[b].max() < max(int(c[b]), 1) and p[b].min() >= (-1 if c[b] == 0 else 0)
