"""EpipolarProcessor with the reference's interface (epipolar_processor.py:8-265).

The 8-point RANSAC (epipolar_processor.py:140-247) evaluates all its hypotheses on the GPU
(sfmba_fund_ransac): the eight indices of every hypothesis are drawn with
Python's `random.sample` in the reference's order, the library computes every fundamental matrix (8x9 null
vector, rank-2 projection, division by f[2][2]), counts the inliers of |x_r^T F x_l| < threshold over all
pairs and returns the first strict maximum.  Normalisation / de-normalisation of the pairs and the
essential-matrix extraction are O(n) and 3x3 host arithmetic.
"""
from __future__ import annotations

import logging
import random

import numpy as np

from . import _capi
from .engine import get_engine


class EpipolarProcessor:
    def __init__(self, ransac_config, engine=None):
        self.fund_mat = np.identity(3)                 # epipolar_processor.py:13-20
        self.esse_mat = np.identity(3)
        self.ransac = ransac_config
        self._engine = engine

    @property
    def engine(self):
        if self._engine is None:
            self._engine = get_engine()
        return self._engine

    # ------------------------------------------------------------------------------------------
    def determine_fundamental_mat(self, matched_pairs, ransac_config=None):
        """matched_pairs = [ref_pts, que_pts] (KeyPt, 3 x n).  Sets self.fund_mat, returns the inlier indices
        (epipolar_processor.py:23-56)."""
        pairs, t_left, t_right = self._normalize(matched_pairs)
        cfg = self.ransac if ransac_config is None else ransac_config
        inliers, fund = self._estimate_ransac(pairs, cfg)
        f = t_right.transpose() @ fund @ t_left        # epipolar_processor.py:250-265
        self.fund_mat = f / f[2][2]
        return inliers

    # ------------------------------------------------------------------------------------------
    def extract_essential_mat(self, left_intrinsic_mat, right_intrinsic_mat):
        """E = K'^T F K projected onto singular values (1, 1, 0) (epipolar_processor.py:58-93)."""
        esse = np.asarray(right_intrinsic_mat).transpose() @ self.fund_mat @ np.asarray(left_intrinsic_mat)
        u, _, vh = np.linalg.svd(esse)
        esse = u @ np.diag([1, 1, 0]) @ vh
        rank = np.linalg.matrix_rank(esse)
        if rank != 2:
            logging.warning('%s : extracted essential matrix is not with rank=2, %d', self.__class__.__name__, rank)
            raise ValueError('esse_mat rank is not equal to 2 : {}'.format(rank))
        self.esse_mat = esse / esse[2][2]

    # ------------------------------------------------------------------------------------------
    @staticmethod
    def _normalize(matched_pairs):
        """Hartley normalisation of both point sets (epipolar_processor.py:96-137)."""
        left = np.asarray(matched_pairs[0], dtype=float)[0:2, :].T
        right = np.asarray(matched_pairs[1], dtype=float)[0:2, :].T
        n = left.shape[0]
        lm = np.average(left, axis=0)
        rm = np.average(right, axis=0)
        ls = (2 * n) ** 0.5 / np.sum(np.sum((left - lm) ** 2, axis=1) ** 0.5)
        rs = (2 * n) ** 0.5 / np.sum(np.sum((right - rm) ** 2, axis=1) ** 0.5)
        t_left = np.array([[ls, 0, -lm[0] * ls], [0, ls, -lm[1] * ls], [0, 0, 1]])
        t_right = np.array([[rs, 0, -rm[0] * rs], [0, rs, -rm[1] * rs], [0, 0, 1]])
        one = np.ones((n, 1))
        ln = (t_left @ np.column_stack((left, one)).T).T
        rn = (t_right @ np.column_stack((right, one)).T).T
        return np.column_stack((ln[:, 0:2], rn[:, 0:2])), t_left, t_right

    # ------------------------------------------------------------------------------------------
    def _estimate_ransac(self, pairs, ransac_config):
        """epipolar_processor.py:194-247: < 8 pairs raises, exactly 8 are used directly, otherwise RANSAC."""
        rows = pairs.shape[0]
        if rows < 8:
            logging.error('%s : number of matched pairs needs equal or more than eight', self.__class__.__name__)
            raise ValueError('Insufficient matched pairs : {}'.format(rows))
        if rows == 8:
            samples = np.arange(8, dtype=np.int32).reshape(1, 8)
            res = self.engine.fund_ransac(pairs, samples, np.inf, want_all=True)
            self._check_rank(res)
            return [0, 1, 2, 3, 4, 5, 6, 7], res['models'][0].reshape(3, 3)
        samples = np.array([random.sample(range(rows), 8) for _ in range(ransac_config.iteration)],
                           dtype=np.int32).reshape(-1, 8)
        res = self.engine.fund_ransac(pairs, samples, ransac_config.inlier_threshold)
        self._check_rank(res)
        if res['hyp'] < 0:
            return None, np.zeros((3, 3))              # no hypothesis had an inlier (epipolar_processor.py:219-220)
        return [int(i) for i in np.flatnonzero(res['mask'])], res['F']

    def _check_rank(self, res):
        if res['flags'] & _capi.FLAG_SINGULAR:         # epipolar_processor.py:179-183
            logging.warning('%s : estimated fundamental matrix is not with rank=2', self.__class__.__name__)
            raise ValueError('f__ rank is not equal to 2')
