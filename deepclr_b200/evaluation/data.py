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


def _vec_to_mat(v) -> np.ndarray:
    """12 row-major numbers -> homogeneous 4x4."""
    return np.vstack([np.asarray(v, dtype=np.float64).reshape(3, 4), [0.0, 0.0, 0.0, 1.0]])


def _mat_to_vec(m) -> np.ndarray:
    return np.asarray(m)[:3, :4].reshape(12)


def _load_rows(filename: str) -> np.ndarray:
    return np.loadtxt(filename, ndmin=2)


class Motion:
    """`transforms[i]` moves frame i to frame i+1, `poses[i]` is the accumulated pose of frame i (poses[0] = identity when
    fed with transforms), `distances[i]` the path length travelled up to frame i."""

    def __init__(self) -> None:
        self.transforms: List[np.ndarray] = []
        self.poses: List[np.ndarray] = []
        self.distances: List[float] = []

    def _advance(self, step: np.ndarray, pose: np.ndarray) -> None:
        travelled = self.distances[-1] + float(np.linalg.norm(step[:3, 3]))
        self.transforms += [step]
        self.poses += [pose]
        self.distances += [travelled]

    def add_transform(self, m: np.ndarray) -> None:
        """Relative motion of the next frame; the first call also creates the identity start pose."""
        if not self.poses:
            self.poses, self.distances = [np.eye(4)], [0]
        self._advance(m, self.poses[-1] @ m)

    def add_pose(self, m: np.ndarray) -> None:
        """Absolute pose of the next frame; from the second one on the transform to the previous pose is derived."""
        if self.poses:
            self._advance(np.linalg.inv(self.poses[-1]) @ m, m)
        else:
            self.poses, self.distances = [m], [0]

    def get_path(self) -> np.ndarray:
        return np.stack([p[:3, 3] for p in self.poses]) if self.poses else np.zeros((0, 3))

    def get_frame_by_distance(self, first_frame: int, distance: float) -> int:
        """First frame that lies more than `distance` of travelled path behind `first_frame`, -1 if none."""
        d = np.asarray(self.distances[first_frame:], dtype=np.float64)
        hits = np.nonzero(d > self.distances[first_frame] + distance)[0]
        return int(first_frame + hits[0]) if hits.size else -1

    @classmethod
    def read(cls, filename: str, has_poses: bool) -> Motion:
        """As the reference (data.py:63-72): `has_poses=True` treats the rows as relative transforms."""
        out = cls()
        feed = out.add_transform if has_poses else out.add_pose
        for row in _load_rows(filename):
            feed(_vec_to_mat(row[:12]))
        return out

    def write(self, filename: str, use_poses: bool) -> None:
        np.savetxt(filename, np.array([_mat_to_vec(m) for m in (self.poses if use_poses else self.transforms)]))


class Sequence:
    """Predicted and ground-truth motion of one run with the frame stamps and inference times."""

    def __init__(self) -> None:
        self.prediction, self.ground_truth = Motion(), Motion()
        self.stamps: list = []
        self.times: list = []

    def _push(self, how: str, stamp: float, pred: np.ndarray, gt: np.ndarray, time: float) -> None:
        self.stamps += [stamp]
        self.times += [time]
        getattr(self.prediction, how)(pred)
        getattr(self.ground_truth, how)(gt)

    def add_transforms(self, stamp: float, pred: np.ndarray, gt: np.ndarray, time: float = 0) -> None:
        self._push('add_transform', stamp, pred, gt, time)

    def add_poses(self, stamp: float, pred: np.ndarray, gt: np.ndarray, time: float = 0) -> None:
        self._push('add_pose', stamp, pred, gt, time)

    @classmethod
    def read(cls, filename: str) -> Sequence:
        out = cls()
        for row in _load_rows(filename):
            out.add_transforms(row[0], _vec_to_mat(row[1:13]), _vec_to_mat(row[13:25]), row[25])
        return out

    @classmethod
    def read_separate(cls, filename_pred: str, filename_gt: str, has_poses: bool) -> Sequence:
        out = cls()
        out.prediction, out.ground_truth = Motion.read(filename_pred, has_poses), Motion.read(filename_gt, has_poses)
        n = len(out.prediction.transforms)
        if n != len(out.ground_truth.transforms):
            raise RuntimeError("Sizes of prediction and ground truth files do not match.")
        out.stamps, out.times = list(range(n)), [0.0] * n
        return out

    def write(self, filename: str) -> None:
        # one row per TRANSFORM: fed with poses there is one transform less than stamps, the surplus stamp is dropped
        rows = zip(self.stamps, self.prediction.transforms, self.ground_truth.transforms, self.times)
        table = [np.concatenate(([s], _mat_to_vec(p), _mat_to_vec(g), [t])) for s, p, g, t in rows]
        np.savetxt(filename, np.array(table))
