"""LabelType: pose label conventions of the reference (deepclr/data/labels.py:10-101), with the transforms3d calls
it relies on restated in numpy (transforms3d is not installed here)."""
from enum import auto
from typing import List, Optional

import numpy as np

from ..config.config import ConfigEnum


def _quat2mat(q: np.ndarray) -> np.ndarray:
    w, x, y, z = (float(v) for v in q)
    n = w * w + x * x + y * y + z * z
    if n < np.finfo(np.float64).eps:
        return np.eye(3)
    s = 2.0 / n
    return np.array([[1 - s * (y * y + z * z), s * (x * y - w * z), s * (x * z + w * y)],
                     [s * (x * y + w * z), 1 - s * (x * x + z * z), s * (y * z - w * x)],
                     [s * (x * z - w * y), s * (y * z + w * x), 1 - s * (x * x + y * y)]])


def _mat2quat(m: np.ndarray) -> np.ndarray:
    # Bar-Itzhack eigen decomposition, the method transforms3d.quaternions.mat2quat uses
    qxx, qyx, qzx, qxy, qyy, qzy, qxz, qyz, qzz = m[:3, :3].flat
    k = np.array([[qxx - qyy - qzz, 0, 0, 0], [qyx + qxy, qyy - qxx - qzz, 0, 0],
                  [qzx + qxz, qzy + qyz, qzz - qxx - qyy, 0],
                  [qyz - qzy, qzx - qxz, qxy - qyx, qxx + qyy + qzz]]) / 3.0
    vals, vecs = np.linalg.eigh(k)
    q = vecs[[3, 0, 1, 2], np.argmax(vals)]
    return -q if q[0] < 0 else q


def _qmult(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    w1, x1, y1, z1 = a
    w2, x2, y2, z2 = b
    return np.array([w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                     w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2, w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2])


def _euler2mat(ai: float, aj: float, ak: float) -> np.ndarray:
    # static xyz ('sxyz'): R = Rz(ak) Ry(aj) Rx(ai)
    ci, si, cj, sj, ck, sk = np.cos(ai), np.sin(ai), np.cos(aj), np.sin(aj), np.cos(ak), np.sin(ak)
    return np.array([[cj * ck, si * sj * ck - ci * sk, ci * sj * ck + si * sk],
                     [cj * sk, si * sj * sk + ci * ck, ci * sj * sk - si * ck],
                     [-sj, si * cj, ci * cj]])


def _mat2euler(m: np.ndarray):
    cy = np.sqrt(m[0, 0] ** 2 + m[1, 0] ** 2)
    if cy > np.finfo(np.float64).eps * 4.0:
        return np.arctan2(m[2, 1], m[2, 2]), np.arctan2(-m[2, 0], cy), np.arctan2(m[1, 0], m[0, 0])
    return np.arctan2(-m[1, 2], m[1, 1]), np.arctan2(-m[2, 0], cy), 0.0


class LabelType(ConfigEnum):
    POSE3D_EULER = auto()
    POSE3D_QUAT = auto()
    POSE3D_DUAL_QUAT = auto()

    @property
    def dim(self) -> int:
        return {LabelType.POSE3D_EULER: 6, LabelType.POSE3D_QUAT: 7, LabelType.POSE3D_DUAL_QUAT: 8}[self]

    @property
    def names(self) -> List[str]:
        return {LabelType.POSE3D_EULER: ['x', 'y', 'z', 'roll', 'pitch', 'yaw'],
                LabelType.POSE3D_QUAT: ['pos_x', 'pos_y', 'pos_z', 'rot_w', 'rot_x', 'rot_y', 'rot_z'],
                LabelType.POSE3D_DUAL_QUAT: ['real_w', 'real_x', 'real_y', 'real_z',
                                             'dual_w', 'dual_x', 'dual_y', 'dual_z']}[self]

    @property
    def bias(self) -> Optional[List[float]]:
        return {LabelType.POSE3D_EULER: None,
                LabelType.POSE3D_QUAT: [0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0],
                LabelType.POSE3D_DUAL_QUAT: [1.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0]}[self]

    @property
    def native_id(self) -> int:
        """label_type code of dclr_arch (include/deepclr_b200.h)."""
        return {LabelType.POSE3D_EULER: 0, LabelType.POSE3D_QUAT: 1, LabelType.POSE3D_DUAL_QUAT: 2}[self]

    def from_matrix(self, data: np.ndarray, scale: Optional[float] = None) -> np.ndarray:
        t, r = data[:3, 3], data[:3, :3]
        if self == LabelType.POSE3D_EULER:
            label = np.array([*t, *np.rad2deg(_mat2euler(r))])
        elif self == LabelType.POSE3D_QUAT:
            label = np.array([*t, *_mat2quat(r)])
        else:
            real = _mat2quat(r)
            label = np.array([*real, *(0.5 * _qmult(np.array([0.0, *t]), real))])
        return label * scale if scale is not None else label

    def to_matrix(self, label: np.ndarray, scale: Optional[float] = None) -> np.ndarray:
        label = np.asarray(label, dtype=np.float64)
        if scale is not None:
            label = label / scale
        m = np.eye(4)
        if self == LabelType.POSE3D_EULER:
            m[:3, 3] = label[:3]
            m[:3, :3] = _euler2mat(*np.deg2rad(label[3:6]))
        elif self == LabelType.POSE3D_QUAT:
            m[:3, 3] = label[:3]
            m[:3, :3] = _quat2mat(label[3:])
        else:
            nrm = np.sqrt(np.dot(label[:4], label[:4])) + 1e-8       # labels.py:46-51
            real, dual = label[:4] / nrm, label[4:] / nrm
            m[:3, :3] = _quat2mat(real)
            m[:3, 3] = (2.0 * _qmult(dual, real * np.array([1.0, -1.0, -1.0, -1.0])))[1:]
        return m
