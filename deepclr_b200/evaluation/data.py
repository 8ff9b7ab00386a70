"""Result-file formats of the reference's evaluation (deepclr/evaluation/data.py:16-137), so that files written after a
run on this package feed the reference's scripts/evaluation.py and scripts/paper/* unchanged:

  Sequence file: one row per frame = [stamp, prediction 3x4 row-major (12), ground truth 3x4 (12), inference time]
                 (26 columns, numpy.savetxt text)
  Motion file:   one 3x4 row-major matrix per row (12 columns): poses (KITTI devkit format) or relative transforms

`Motion` keeps transforms, accumulated poses and travelled distance in step, whichever of the two it is fed with.
"""
from __future__ import annotations

from typing import List

import numpy as np


def _vec_to_mat(v: np.ndarray) -> np.ndarray:
    m = np.eye(4)
    m[:3, :4] = np.asarray(v, dtype=np.float64).reshape(3, 4)
    return m


def _mat_to_vec(m: np.ndarray) -> np.ndarray:
    return np.asarray(m)[:3, :4].reshape(12)


class Motion:
    def __init__(self) -> None:
        self.transforms: List[np.ndarray] = []
        self.poses: List[np.ndarray] = []
        self.distances: List[float] = []

    def _step(self, transform: np.ndarray) -> None:
        self.transforms.append(transform)
        self.distances.append(self.distances[-1] + float(np.linalg.norm(transform[:3, 3])))

    def add_transform(self, m: np.ndarray) -> None:
        """Relative motion of the next frame; the first call also creates the identity start pose."""
        if not self.poses:
            self.poses.append(np.eye(4))
            self.distances.append(0)
        self.poses.append(self.poses[-1] @ m)
        self._step(m)

    def add_pose(self, m: np.ndarray) -> None:
        """Absolute pose of the next frame; from the second one on the transform to the previous pose is derived."""
        self.poses.append(m)
        if len(self.poses) == 1:
            self.distances.append(0)
        else:
            self._step(np.linalg.inv(self.poses[-2]) @ m)

    def get_path(self) -> np.ndarray:
        return np.array([p[:3, 3] for p in self.poses])

    def get_frame_by_distance(self, first_frame: int, distance: float) -> int:
        """First frame that lies more than `distance` of travelled path behind `first_frame`, -1 if none."""
        limit = self.distances[first_frame] + distance
        for i in range(first_frame, len(self.distances)):
            if self.distances[i] > limit:
                return i
        return -1

    @classmethod
    def read(cls, filename: str, has_poses: bool) -> Motion:
        """As the reference (data.py:63-72): `has_poses=True` treats the rows as relative transforms."""
        motion = cls()
        for row in np.loadtxt(filename, ndmin=2):
            (motion.add_transform if has_poses else motion.add_pose)(_vec_to_mat(row[:12]))
        return motion

    def write(self, filename: str, use_poses: bool) -> None:
        np.savetxt(filename, np.array([_mat_to_vec(m) for m in (self.poses if use_poses else self.transforms)]))


class Sequence:
    def __init__(self) -> None:
        self.prediction = Motion()
        self.ground_truth = Motion()
        self.stamps: list = []
        self.times: list = []

    def add_transforms(self, stamp: float, pred: np.ndarray, gt: np.ndarray, time: float = 0) -> None:
        self.stamps.append(stamp)
        self.times.append(time)
        self.prediction.add_transform(pred)
        self.ground_truth.add_transform(gt)

    def add_poses(self, stamp: float, pred: np.ndarray, gt: np.ndarray, time: float = 0) -> None:
        self.stamps.append(stamp)
        self.times.append(time)
        self.prediction.add_pose(pred)
        self.ground_truth.add_pose(gt)

    @classmethod
    def read(cls, filename: str) -> Sequence:
        seq = cls()
        for row in np.loadtxt(filename, ndmin=2):
            seq.add_transforms(row[0], _vec_to_mat(row[1:13]), _vec_to_mat(row[13:25]), row[25])
        return seq

    @classmethod
    def read_separate(cls, filename_pred: str, filename_gt: str, has_poses: bool) -> Sequence:
        seq = cls()
        seq.prediction = Motion.read(filename_pred, has_poses)
        seq.ground_truth = Motion.read(filename_gt, has_poses)
        n = len(seq.prediction.transforms)
        if len(seq.ground_truth.transforms) != n:
            raise RuntimeError("Sizes of prediction and ground truth files do not match.")
        seq.stamps = np.arange(n).tolist()
        seq.times = np.zeros(n).tolist()
        return seq

    def write(self, filename: str) -> None:
        rows = [np.concatenate(([s], _mat_to_vec(p), _mat_to_vec(g), [t])) for s, p, g, t in
                zip(self.stamps, self.prediction.transforms, self.ground_truth.transforms, self.times)]
        np.savetxt(filename, np.array(rows))
