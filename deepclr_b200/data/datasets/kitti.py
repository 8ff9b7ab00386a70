"""Dependency-free KITTI odometry reader with the interface of the reference's
deepclr/data/datasets/kitti.py:27-84 (KittiOdometryVelodyneData / ...SequenceData yield the same dicts), minus its
third-party layers: `pykitti.odometry` (file layout, calib.txt `Tr:` row, times.txt, poses/<seq>.txt) is restated in
`KittiOdometry`, `dataflow.RNGDataFlow` is replaced by a plain iterable with its own numpy generator.

On-disk formats:
  sequences/<seq>/velodyne/<000000>.bin   float32 little endian, rows of (x, y, z, reflectance)
  sequences/<seq>/calib.txt               `P0: .. P3:` 3x4 projections, `Tr:` 3x4 velodyne -> cam0
  sequences/<seq>/times.txt               one timestamp [s] per frame
  poses/<seq>.txt                         one 3x4 cam0 pose per frame (absent for the test sequences 11-21)

`stage_scans` reads raw scans straight into ONE pinned staging buffer (no per-scan tensors, no concatenate), ready for
a single H2D copy into a `RaggedClouds` batch -> dclr_filter_clouds -> dclr_subsample_clouds -> forward.
"""
import glob
import os
from datetime import timedelta
from types import SimpleNamespace
from typing import Iterator, List, Optional, Sequence, Tuple

import numpy as np


def _timedelta_to_us(td: timedelta) -> float:
    return (td.days * 86400 + td.seconds) * 1000000 + td.microseconds


def cam2velo(p: np.ndarray, v: np.ndarray) -> np.ndarray:
    """Camera-frame pose -> velodyne-frame pose with the calibration v = T_cam0_velo (kitti.py:15-18)."""
    return np.linalg.inv(v) @ p @ v


def velo2cam(p: np.ndarray, v: np.ndarray) -> np.ndarray:
    """Velodyne-frame pose -> camera-frame pose (kitti.py:21-24)."""
    return v @ p @ np.linalg.inv(v)


def read_velodyne_bin(path: str) -> np.ndarray:
    """[n, 4] float32 (x, y, z, reflectance)."""
    scan = np.fromfile(path, dtype='<f4')
    if scan.size % 4:
        raise RuntimeError(f'{path}: size is not a multiple of 4 floats')
    return scan.reshape(-1, 4)


def _as_4x4(row12: np.ndarray) -> np.ndarray:
    m = np.eye(4)
    m[:3, :] = np.asarray(row12, dtype=np.float64).reshape(3, 4)
    return m


class KittiOdometry:
    """The subset of `pykitti.odometry(base_path, sequence)` the reference uses: `len()`, `.timestamps` (timedelta),
    `.poses` (list of 4x4, empty without ground truth), `.calib.T_cam0_velo`, `.velo_files`, `.get_velo(k)`."""

    def __init__(self, base_path: str, sequence: str, frames: Optional[Sequence[int]] = None):
        self.sequence = sequence
        self.sequence_path = os.path.join(base_path, 'sequences', sequence)
        self.pose_path = os.path.join(base_path, 'poses')
        self.velo_files = sorted(glob.glob(os.path.join(self.sequence_path, 'velodyne', '*.bin')))
        if frames is not None:
            self.velo_files = [self.velo_files[i] for i in frames]
        self.calib = self._load_calib()
        self.timestamps = self._load_timestamps(frames)
        self.poses = self._load_poses(frames)

    def __len__(self) -> int:
        return len(self.timestamps)

    def _load_calib(self) -> SimpleNamespace:
        entries = {}
        path = os.path.join(self.sequence_path, 'calib.txt')
        if os.path.exists(path):
            with open(path) as f:
                for line in f:
                    if ':' in line:
                        key, vals = line.split(':', 1)
                        entries[key.strip()] = np.array([float(v) for v in vals.split()])
        calib = SimpleNamespace(**{f'P_rect_{i}0': entries[f'P{i}'].reshape(3, 4) for i in range(4) if f'P{i}' in entries})
        calib.T_cam0_velo = _as_4x4(entries['Tr']) if 'Tr' in entries else np.eye(4)
        return calib

    def _load_timestamps(self, frames) -> List[timedelta]:
        path = os.path.join(self.sequence_path, 'times.txt')
        stamps = []
        if os.path.exists(path):
            with open(path) as f:
                stamps = [timedelta(seconds=float(line)) for line in f if line.strip()]
        if frames is not None:
            stamps = [stamps[i] for i in frames]
        return stamps

    def _load_poses(self, frames) -> List[np.ndarray]:
        path = os.path.join(self.pose_path, self.sequence + '.txt')
        if not os.path.exists(path):
            return []
        rows = np.loadtxt(path, dtype=np.float64, ndmin=2)
        poses = [_as_4x4(r) for r in rows]
        if frames is not None:
            poses = [poses[i] for i in frames]
        return poses

    def get_velo(self, idx: int) -> np.ndarray:
        return read_velodyne_bin(self.velo_files[idx])


class _Flow:
    """What the reference takes from dataflow.RNGDataFlow: reset_state() seeds self.rng; iteration may shuffle."""
    def __init__(self, shuffle: bool, seed: Optional[int] = None):
        self.shuffle = shuffle
        self.rng = np.random.default_rng(seed)

    def reset_state(self, seed: Optional[int] = None) -> None:
        self.rng = np.random.default_rng(seed)


class KittiOdometryVelodyneData(_Flow):
    """Velodyne clouds and velodyne-frame poses of one sequence, frame by frame (kitti.py:27-49): dicts with `idx`,
    `timestamp` [us], `pose` 4x4 (identity without ground truth), `cloud` [n,4] float32."""

    def __init__(self, base_path: str, sequence: str, shuffle: bool = False, seed: Optional[int] = None):
        super().__init__(shuffle, seed)
        self.data = KittiOdometry(base_path, sequence)
        self.calib = self.data.calib.T_cam0_velo

    def __len__(self) -> int:
        return len(self.data)

    def _frame(self, k: int) -> dict:
        pose = np.eye(4) if len(self.data.poses) == 0 else cam2velo(self.data.poses[k], self.calib)
        return {'idx': k, 'timestamp': _timedelta_to_us(self.data.timestamps[k]), 'pose': pose,
                'cloud': self.data.get_velo(k)}

    def __iter__(self) -> Iterator[dict]:
        order = list(range(len(self.data)))
        if self.shuffle:
            self.rng.shuffle(order)
        for k in order:
            yield self._frame(k)


class KittiOdometryVelodyneSequenceData(KittiOdometryVelodyneData):
    """Windows of `seq_length` consecutive frames every `seq_step` frames (kitti.py:52-96): lists of frame dicts."""

    def __init__(self, base_path: str, sequence: str, seq_length: int, seq_step: int = 1, shuffle: bool = False,
                 seed: Optional[int] = None):
        super().__init__(base_path, sequence, shuffle, seed)
        self.seq_length, self.seq_step = int(seq_length), int(seq_step)
        assert self.seq_length > 0 and self.seq_step > 0
        self.idxs = list(range(0, len(self.data) - self.seq_length + 1, self.seq_step))

    def __len__(self) -> int:
        return len(self.idxs)

    def __iter__(self) -> Iterator[List[dict]]:
        starts = list(self.idxs)
        if self.shuffle:
            self.rng.shuffle(starts)
        for s in starts:
            yield [self._frame(k) for k in range(s, s + self.seq_length)]


def stage_scans(files: Sequence[str], pinned: bool = True) -> Tuple['torch.Tensor', List[int]]:  # noqa: F821
    """Read raw velodyne .bin files back to back into one (pinned) [total, 4] float32 staging tensor with readinto --
    no intermediate arrays.  Returns (stage, lengths); `RaggedClouds.from_staged(stage, lengths, device)` uploads it."""
    import torch
    sizes = [os.path.getsize(f) for f in files]
    if any(s % 16 for s in sizes):
        raise RuntimeError('velodyne file size is not a multiple of 16 bytes')
    lengths = [s // 16 for s in sizes]
    stage = torch.empty((sum(lengths), 4), dtype=torch.float32, pin_memory=pinned and torch.cuda.is_available())
    flat = stage.numpy().reshape(-1).view(np.uint8)
    pos = 0
    for f, s in zip(files, sizes):
        with open(f, 'rb', buffering=0) as fh:
            got = fh.readinto(memoryview(flat[pos:pos + s]))
        if got != s:
            raise RuntimeError(f'{f}: short read ({got} of {s} bytes)')
        pos += s
    return stage, lengths
