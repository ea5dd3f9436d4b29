"""Math utilities and small types with the reference's names and semantics (utils.py:6-174).

Written from the behaviour, not copied: KeyPt/TriPt are homogeneous column-per-point float64
arrays; the quaternion helpers keep the reference's conventions ((w,x,y,z), qw > 0 branch only,
ValueError on invalid input)."""
from __future__ import annotations

import math
import random

import numpy as np


class KeyPt(np.ndarray):
    """2D points, homogeneous, one column per point: shape (3, n), last row 1 (utils.py:6-13)."""

    def __new__(cls, num):
        arr = np.zeros((3, num), dtype=float)
        arr[2, :] = 1.0
        return arr.view(cls)


class TriPt(np.ndarray):
    """3D points, homogeneous, one column per point: shape (4, n), last row 1 (utils.py:17-24)."""

    def __new__(cls, num):
        arr = np.zeros((4, num), dtype=float)
        arr[3, :] = 1.0
        return arr.view(cls)


def verify_rotation_mat(rot) -> bool:
    """One-sided 1e-8 test on det(R) - 1 and on inv(R) - R^T (utils.py:101-105)."""
    rot = np.asarray(rot)
    if rot.shape != (3, 3):
        return False
    if np.linalg.det(rot) - 1 >= 1e-8:
        return False
    return not np.any((np.linalg.inv(rot) - rot.T) > 1e-8)


def convert_rotation_to_quaternion(rot_mat) -> np.ndarray:
    """R (3,3) -> q (4,1) in (w,x,y,z) order; raises ValueError for an invalid R or |qw| < 1e-6
    (utils.py:28-60)."""
    if not verify_rotation_mat(rot_mat):
        raise ValueError('{} : Invalid input rotation matrix \n {}'.format('convert_rotation_to_quaternion', rot_mat))
    r = np.asarray(rot_mat, dtype=float)
    qw = math.sqrt(1 + r[0][0] + r[1][1] + r[2][2]) / 2.0
    if abs(qw) < 1e-6:
        raise ValueError('{} : Invalid output qw \n {}'.format('convert_rotation_to_quaternion', qw))
    s = 4 * qw
    return np.array([[qw], [(r[2][1] - r[1][2]) / s], [(r[0][2] - r[2][0]) / s], [(r[1][0] - r[0][1]) / s]])


def quaternion_to_rotation_unchecked(q) -> np.ndarray:
    w, x, y, z = (float(v) for v in np.asarray(q, dtype=float).reshape(4))
    return np.array([
        [1 - 2 * z * z - 2 * y * y, -2 * z * w + 2 * y * x, 2 * y * w + 2 * z * x],
        [2 * x * y + 2 * w * z, 1 - 2 * z * z - 2 * x * x, 2 * z * y - 2 * x * w],
        [2 * x * z - 2 * w * y, 2 * y * z + 2 * w * x, 1 - 2 * y * y - 2 * x * x]])


def convert_quaternion_to_rotation(quaternion) -> np.ndarray:
    """q -> R (3,3); raises ValueError if the result fails verify_rotation_mat (utils.py:64-97)."""
    rot = quaternion_to_rotation_unchecked(quaternion)
    if not verify_rotation_mat(rot):
        raise ValueError('{} : Invalid output rotation matrix \n {}'.format('convert_quaternion_to_rotation', rot))
    return rot


class LMConfig:
    """Levenberg-Marquardt configuration (utils.py:109-125; unused by the reference's solvers)."""

    def __init__(self, stop_ratio, iteration):
        self.stop_ratio = stop_ratio
        self.iteration = iteration


class RansacConfig:
    """RANSAC parameters (utils.py:129-174).  Like the reference, construction reseeds the global
    `random` module with -1 unless is_use_seed is False, and raises the iteration count to the
    value the confidences imply."""

    def __init__(self, inlier_threshold, subset_confidence, sample_confidence, sample_num, iteration,
                 is_use_seed=True):
        self.inlier_threshold = inlier_threshold
        self.subset_confidence = subset_confidence
        self.sample_confidence = sample_confidence
        self.sample_num = int(sample_num)
        self.iteration = int(iteration)
        self.random_seed = -1
        needed = math.log(1.0 - subset_confidence) / math.log(1.0 - math.pow(sample_confidence, sample_num))
        if needed > iteration:
            print('RANSAC : iteration increases from {} to {}'.format(iteration, needed))
            self.iteration = int(needed)
        if is_use_seed:
            random.seed(self.random_seed)
