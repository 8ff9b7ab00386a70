"""CPU suite for the data formats either side of the path: the dependency-free KITTI odometry
reader and the result files the reference's evaluation scripts read."""
import importlib
import os
import sys

import numpy as np
import pytest

from deepclr_b200.data.datasets import (KittiOdometryVelodyneData, KittiOdometryVelodyneSequenceData, cam2velo,
                                        stage_scans, velo2cam)
from deepclr_b200.evaluation import Motion, Sequence


def _pose(rng):
    q = rng.normal(size=4)
    q /= np.linalg.norm(q)
    w, x, y, z = q
    m = np.eye(4)
    m[:3, :3] = [[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                 [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                 [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]]
    m[:3, 3] = rng.normal(size=3)
    return m


@pytest.fixture()
def kitti_tree(tmp_path):
    rng = np.random.default_rng(0)
    seq = tmp_path / 'sequences' / '04'
    (seq / 'velodyne').mkdir(parents=True)
    (tmp_path / 'poses').mkdir()
    scans = [rng.normal(size=(n, 4)).astype('<f4') for n in (100, 57, 0, 311, 64)]
    for i, s in enumerate(scans):
        s.tofile(seq / 'velodyne' / f'{i:06d}.bin')
    tr = _pose(rng)
    with open(seq / 'calib.txt', 'w') as f:
        for i in range(4):
            f.write(f'P{i}: ' + ' '.join(f'{v:.12e}' for v in rng.normal(size=12)) + '\n')
        f.write('Tr: ' + ' '.join(f'{v:.12e}' for v in tr[:3].reshape(-1)) + '\n')
    times = [0.0, 0.103, 0.2071, 0.31, 1.5]
    (seq / 'times.txt').write_text(''.join(f'{t:.6e}\n' for t in times))
    poses = [_pose(rng) for _ in scans]
    np.savetxt(tmp_path / 'poses' / '04.txt', np.array([p[:3].reshape(-1) for p in poses]))
    return str(tmp_path), scans, tr, times, poses


def test_kitti_reader_yields_reference_dicts(kitti_tree):
    base, scans, tr, times, poses = kitti_tree
    ds = KittiOdometryVelodyneData(base, '04')
    assert len(ds) == 5
    frames = list(ds)
    for k, fr in enumerate(frames):
        assert fr['idx'] == k and fr['cloud'].dtype == np.float32 and np.array_equal(fr['cloud'], scans[k])
        assert fr['timestamp'] == pytest.approx(times[k] * 1e6, abs=1)
        np.testing.assert_allclose(fr['pose'], np.linalg.inv(tr) @ poses[k] @ tr, atol=1e-9)
        np.testing.assert_allclose(velo2cam(fr['pose'], ds.calib), poses[k], atol=1e-9)
    np.testing.assert_allclose(cam2velo(poses[1], tr), frames[1]['pose'])
    sh = KittiOdometryVelodyneData(base, '04', shuffle=True, seed=3)
    assert sorted(f['idx'] for f in sh) == [0, 1, 2, 3, 4]
    win = KittiOdometryVelodyneSequenceData(base, '04', seq_length=2, seq_step=2)
    assert len(win) == 2 and [[f['idx'] for f in w] for w in win] == [[0, 1], [2, 3]]
    os.remove(os.path.join(base, 'poses', '04.txt'))            # test sequences: identity poses
    assert all(np.array_equal(f['pose'], np.eye(4)) for f in KittiOdometryVelodyneData(base, '04'))


def test_stage_scans_is_the_concatenation(kitti_tree):
    base, scans, *_ = kitti_tree
    files = KittiOdometryVelodyneData(base, '04').data.velo_files
    stage, lengths = stage_scans(files, pinned=False)
    assert lengths == [s.shape[0] for s in scans]
    assert np.array_equal(stage.numpy(), np.concatenate(scans))
    with open(files[1], 'ab') as f:
        f.write(b'\0' * 5)
    with pytest.raises(RuntimeError):
        stage_scans(files, pinned=False)


def test_result_files_round_trip_and_layout(tmp_path):
    rng = np.random.default_rng(1)
    seq = Sequence()
    preds, gts = [_pose(rng) for _ in range(6)], [_pose(rng) for _ in range(6)]
    for i, (p, g) in enumerate(zip(preds, gts)):
        seq.add_transforms(100.0 + i, p, g, 0.01 * i)
    path = str(tmp_path / 'seq.txt')
    seq.write(path)
    raw = np.loadtxt(path)
    assert raw.shape == (6, 26)
    np.testing.assert_allclose(raw[2, 1:13], preds[2][:3].reshape(-1))
    np.testing.assert_allclose(raw[2, 13:25], gts[2][:3].reshape(-1))
    assert raw[2, 0] == 102.0 and raw[2, 25] == pytest.approx(0.02)
    back = Sequence.read(path)
    assert len(back.prediction.poses) == 7 and len(back.prediction.transforms) == 6
    np.testing.assert_allclose(back.prediction.poses[3], preds[0] @ preds[1] @ preds[2], atol=1e-12)
    assert back.ground_truth.distances[-1] == pytest.approx(sum(np.linalg.norm(g[:3, 3]) for g in gts))
    # poses in, transforms out
    m = Motion()
    for p in back.prediction.poses:
        m.add_pose(p)
    for a, b in zip(m.transforms, preds):
        np.testing.assert_allclose(a, b, atol=1e-9)
    assert m.get_frame_by_distance(0, 1e9) == -1 and m.get_frame_by_distance(0, 0.0) == 1
    assert m.get_path().shape == (7, 3)
    m.write(str(tmp_path / 'poses.txt'), use_poses=True)
    assert np.loadtxt(str(tmp_path / 'poses.txt')).shape == (7, 12)
    m.write(str(tmp_path / 'tf.txt'), use_poses=False)
    both = Sequence.read_separate(str(tmp_path / 'tf.txt'), str(tmp_path / 'tf.txt'), has_poses=True)
    assert len(both.stamps) == 6 and both.times == [0.0] * 6
    np.testing.assert_allclose(both.prediction.poses[-1], m.poses[-1], atol=1e-9)


@pytest.mark.skipif(not os.path.isdir('/root/reference/deepclr/evaluation'), reason='reference tree not mounted')
def test_result_files_equal_the_reference_writer(tmp_path):
    """Build container only: the same sequence written by the reference's own deepclr/evaluation/data.py."""
    spec = importlib.util.spec_from_file_location('ref_eval_data', '/root/reference/deepclr/evaluation/data.py')
    ref = importlib.util.module_from_spec(spec)
    sys.modules['ref_eval_data'] = ref
    spec.loader.exec_module(ref)
    rng = np.random.default_rng(2)
    mine, theirs = Sequence(), ref.Sequence()
    for i in range(5):
        p, g = _pose(rng), _pose(rng)
        mine.add_poses(i, p, g, 0.5 * i)
        theirs.add_poses(i, p, g, 0.5 * i)
    mine.write(str(tmp_path / 'a.txt'))
    theirs.write(str(tmp_path / 'b.txt'))
    assert open(tmp_path / 'a.txt').read() == open(tmp_path / 'b.txt').read()
    for a, b in zip(mine.prediction.distances, theirs.prediction.distances):
        assert a == pytest.approx(b, abs=1e-12)
    mine.prediction.write(str(tmp_path / 'pa.txt'), use_poses=True)
    theirs.prediction.write(str(tmp_path / 'pb.txt'), use_poses=True)
    assert open(tmp_path / 'pa.txt').read() == open(tmp_path / 'pb.txt').read()
    r1, r2 = Sequence.read(str(tmp_path / 'b.txt')), ref.Sequence.read(str(tmp_path / 'a.txt'))
    for a, b in zip(r1.ground_truth.poses, r2.ground_truth.poses):
        np.testing.assert_allclose(a, b, atol=1e-12)


def test_reference_import_names_resolve():
    """`deepclr_b200.install_as_deepclr()` makes the reference's import lines for these modules work unchanged
    (scripts/inference.py, scripts/evaluation.py)."""
    import subprocess
    code = ("import deepclr_b200; deepclr_b200.install_as_deepclr();"
            "from deepclr.evaluation.data import Sequence, Motion;"
            "from deepclr.data.datasets.kitti import KittiOdometryVelodyneData, KittiOdometryVelodyneSequenceData, cam2velo, velo2cam;"
            "from deepclr.data.labels import LabelType;"
            "from deepclr.models import ModelInferenceHelper, build_model, load_trained_model;"
            "from deepclr.config import load_model_config")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    subprocess.check_call([sys.executable, '-c', code], cwd=root)
