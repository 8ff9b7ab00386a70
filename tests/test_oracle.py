"""CPU suite: the oracle against the committed golden vectors (generated from the reference's own Python modules by
oracle/gen_golden.py) and against independent implementations (float64 numpy FPS as in the reference's
deepclr/data/transforms/transforms.py:31-59, scipy cKDTree for ball membership / kNN sets)."""
import numpy as np
import pytest
import torch
from scipy.spatial import cKDTree

from oracle import model_oracle, ops
from tests.util import build_from_golden, golden_input, load_golden


def test_ops_golden():
    g = load_golden('ops_small')
    xyz, small, feats = g['xyz'], g['small'], g['feats']
    assert np.array_equal(ops.fps(xyz, 64), g['fps_ref'])
    assert np.array_equal(ops.fps_lowidx(xyz, 64), g['fps_low'])
    # reference test shape: more samples than points (tests/model/test_deepclr.py:10-25) -> index 0 repeats
    assert np.array_equal(ops.fps(small, 128), g['small_fps_ref'])
    assert np.array_equal(ops.fps_lowidx(small, 128), g['small_fps_low'])
    assert (g['small_fps_low'][:, 96:] == 0).all()
    idx, cnt = ops.ball_query(0.35, 16, xyz, g['new_xyz'], return_cnt=True)
    assert np.array_equal(idx, g['ball_idx']) and np.array_equal(cnt, g['ball_cnt'])
    assert np.array_equal(ops.gather(feats, g['fps_low']), g['gathered'])
    assert np.array_equal(ops.group(feats, idx), g['grouped'])
    assert np.array_equal(ops.knn(xyz[:2].reshape(-1, 3), g['new_xyz'][:2].reshape(-1, 3), 9, 2), g['knn'])


def test_fps_against_float64_numpy():
    rng = np.random.default_rng(1)
    pts = rng.normal(size=(1, 500, 3)).astype(np.float32)
    idx = ops.fps_lowidx(pts, 50)[0]
    p = pts[0].astype(np.float64)
    d = np.full(500, np.inf)
    cur, ref = 0, [0]
    for _ in range(49):
        d = np.minimum(d, ((p - p[cur]) ** 2).sum(1))
        cur = int(np.argmax(d))
        ref.append(cur)
    assert list(idx) == ref          # generic float data: no ties, fp32 and fp64 orders coincide


def test_fps_tie_rules():
    # two coincident points make every later maximum unique or tied at distance 0 only once all points are used
    pts = np.zeros((1, 8, 3), np.float32)
    pts[0, :, 0] = [0, 1, 1, 2, 2, 3, 3, 4]
    idx, ties = ops.fps_lowidx(pts, 8, return_ties=True)
    assert idx[0, 0] == 0 and idx[0, 1] == 7
    assert ties[0] > 0
    assert sorted(set(idx[0].tolist())) != []      # runs through; lowest index wins the exact ties
    assert idx[0, 2] in (3, 4) and idx[0, 2] == 3  # |2-0| = |2-4|: points 3 and 4 tie -> lowest index 3


def test_ball_query_semantics():
    rng = np.random.default_rng(2)
    xyz = rng.uniform(-1, 1, (2, 400, 3)).astype(np.float32)
    centres = xyz[:, :40].copy()
    centres[0, 0] = 50.0                       # empty ball
    r, ns = 0.3, 8
    idx, cnt = ops.ball_query(r, ns, xyz, centres, return_cnt=True)
    assert cnt[0, 0] == 0 and (idx[0, 0] == 0).all()
    for b in range(2):
        tree = cKDTree(xyz[b].astype(np.float64))
        for i in range(40):
            members = sorted(tree.query_ball_point(centres[b, i].astype(np.float64), r - 1e-6))
            got = idx[b, i, :cnt[b, i]].tolist()
            assert got == sorted(got) and len(set(got)) == len(got)
            if len(members) <= ns:
                # boundary points may differ by rounding; everything strictly inside must be there
                assert set(members) <= set(got)
            assert (idx[b, i, cnt[b, i]:] == (got[0] if got else 0)).all()


def test_knn_against_kdtree():
    rng = np.random.default_rng(3)
    x = rng.normal(size=(2 * 200, 3)).astype(np.float32)
    y = rng.normal(size=(2 * 50, 3)).astype(np.float32)
    out = ops.knn(x, y, 7, 2).reshape(2, 100, 7)
    assert (out[0] == np.arange(100)[:, None]).all()
    for b in range(2):
        _, nn = cKDTree(x[b * 200:(b + 1) * 200].astype(np.float64)).query(y[b * 50:(b + 1) * 50].astype(np.float64), 7)
        assert np.array_equal(out[1, b * 50:(b + 1) * 50] - b * 200, nn)
    with pytest.raises(ValueError):
        ops.knn(x[:5], y[:5], 6, 1)


@pytest.mark.parametrize('case', ['modelnet_b2', 'kitti_small_b2', 'kitti_b1'])
def test_model_oracle_reproduces_reference_goldens(case):
    g = load_golden(case)
    model, cfg = build_from_golden(g)
    x = torch.from_numpy(golden_input(g))
    dbg = {}
    y = model_oracle.forward(model.state_dict(), cfg, x, debug=dbg).numpy()
    assert np.array_equal(dbg['fps_idx'], g['fps_idx'])
    assert np.array_equal(dbg['ball_cnt0'], g['ball_cnt0']) and np.array_equal(dbg['ball_cnt1'], g['ball_cnt1'])
    assert np.array_equal(dbg['knn_idx'], g['knn_idx'])
    np.testing.assert_allclose(dbg['cloud_feat'], g['cloud_feat'], rtol=0, atol=1e-5)
    np.testing.assert_allclose(dbg['motion_feat'], g['motion_feat'], rtol=0, atol=1e-5)
    np.testing.assert_allclose(y, g['y'], rtol=0, atol=1e-5)
    assert bool(g['fps_rules_agree'])          # reference block-tree tie-break == lowest-index on these inputs
    rot, trans = model_oracle.pose_errors(y[0], g['y'][0])
    assert rot < 1e-3 and trans < 1e-5


def test_preproc_restatement_matches_reference_transform_goldens():
    """tests/golden/preproc_small.npz holds the outputs of the reference's own SystematicErasing -> RangeSelection ->
    TruncateDimension classes (oracle/gen_golden.py:preproc_case)."""
    from oracle import preproc
    g = load_golden('preproc_small')
    for ci in range(int(g['num_cfgs'])):
        nth, start, rmin, rmax, dim = g[f'cfg_{ci}']
        for b in range(int(g['num_clouds'])):
            got = preproc.filter_cloud(g[f'cloud_{b}'], int(nth), int(start), float(rmin), float(rmax), int(dim))
            assert np.array_equal(got, g[f'out_{ci}_{b}'], equal_nan=True), (ci, b)
    out = g['out_1_0']
    assert not np.isnan(out[:, :2]).any()                       # the NaN row dropped
    assert (out[:, :2] == np.float32([2.0, -1.0])).all(1).any() and (out[:, :2] == np.float32([-50.0, 3.0])).all(1).any()


def test_feistel_permutation_is_a_bijection():
    from oracle import preproc
    for n in (1, 2, 3, 5, 16, 17, 255, 256, 257, 1000, 4097):
        for b in (0, 3):
            key = preproc.cloud_key(99, b)
            img = sorted(preproc.feistel_perm(j, n, key) for j in range(n))
            assert img == list(range(n)), n
    a = preproc.subsample_rows(1000, 100, 1, 0)
    assert len(set(a.tolist())) == 100 and not np.array_equal(a, preproc.subsample_rows(1000, 100, 2, 0))
    assert np.array_equal(preproc.subsample_rows(100, 100, 1, 0), np.arange(100))
    assert np.array_equal(preproc.subsample_rows(0, 4, 1, 0), np.full(4, -1))
