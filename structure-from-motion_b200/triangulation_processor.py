"""TriangulationProcessor with the reference's interface (triangulation_processor.py:7-309); the
nonlinear refinement runs on the GPU (sfmba_triangulate_refine_batched), batched over all points.

Signatures, argument meaning, defaults ("falsy kwargs fall back to the instance value") and error
behaviour follow the reference; the iterative maths lives in csrc/refine.cu.
"""
from __future__ import annotations

import logging

import numpy as np

from .engine import get_engine
from .utils import TriPt


class TriangulationProcessor:
    """Determine 3D positions from matched 2D key points of n views."""

    def __init__(self, damping_factor=0.5, iteration=100, engine=None):
        self.damping_factor = damping_factor        # triangulation_processor.py:12-18
        self.iteration = iteration
        self.tri_pts = None
        self._engine = engine

    @property
    def engine(self):
        if self._engine is None:
            self._engine = get_engine()
        return self._engine

    # ------------------------------------------------------------------------------------------
    def add_tri_pt(self, tri_pt):
        """Append triangulated points (triangulation_processor.py:21-28)."""
        if not np.any(self.tri_pts):
            self.tri_pts = tri_pt
        else:
            self.tri_pts = np.hstack((self.tri_pts, tri_pt))

    # ------------------------------------------------------------------------------------------
    def triangulate(self, projs, matched_pairs, damping_factor=None, iteration=None):
        """Linear then nonlinear triangulation (triangulation_processor.py:31-88)."""
        if not damping_factor:
            damping_factor = self.damping_factor
        if not iteration:
            iteration = self.iteration
        num_projs = len(projs)
        num_views = len(matched_pairs)
        if num_projs != num_views:
            logging.warning('%s : the numbers of views and of projections are different, %d and %d',
                            self.__class__.__name__, num_views, num_projs)
            raise ValueError('different numbers of views and projections : {} - {}'.format(num_views, num_projs))
        if num_projs < 2:
            logging.warning('%s : insufficient projections, %d', self.__class__.__name__, num_projs)
            raise ValueError('insufficient projections : {}'.format(num_projs))
        init_3d_pts = self.linear_triangulate(projs, matched_pairs)
        return self.nonlinear_triangulate(init_3d_pts, projs, matched_pairs, damping_factor, iteration)

    # ------------------------------------------------------------------------------------------
    def linear_triangulate(self, projs, matched_pairs):
        """DLT triangulation, one 2n x 4 SVD per point (triangulation_processor.py:91-157), here one GPU
        thread per point."""
        if len(matched_pairs) != len(projs) != 2:
            print('{}:{} - num of projs {} and matched pairs {} need to be 2'.format(
                self.__class__.__name__, 'linear_triangulate', len(projs), len(matched_pairs)))
            return None
        if matched_pairs[0].shape[1] != matched_pairs[1].shape[1]:
            print('{}:{} - matched pairs number does not match {} vs {}'.format(
                self.__class__.__name__, 'linear_triangulate', matched_pairs[0].shape[1], matched_pairs[1].shape[1]))
            return None
        num_views = len(matched_pairs)
        num_pts = matched_pairs[0].shape[1]
        tri_pts = TriPt(num_pts)
        if num_pts:
            # one thread per point on the GPU (sfmba_triangulate_linear_batched): CSR with num_views entries per point
            rowptr = np.arange(num_pts + 1, dtype=np.int32) * num_views
            obs_view = np.tile(np.arange(num_views, dtype=np.int32), num_pts)
            uv = np.stack([np.asarray(mp, dtype=np.float64)[0:2, :].T for mp in matched_pairs], axis=1)   # (M,n,2)
            pts, _ = self.engine.triangulate_linear(np.stack([np.asarray(p, dtype=np.float64) for p in projs]),
                                                    rowptr, obs_view, uv.reshape(-1, 2))
            tri_pts[0:3, :] = pts.T
            tri_pts[3, :] = 1.0
        return tri_pts

    # ------------------------------------------------------------------------------------------
    def nonlinear_triangulate(self, init_3d_pts, projs, matched_pairs, damping_factor=None, iteration=None):
        """Geometric-error refinement, X -= inv(J^T J + lambda I) J^T (proj - obs) per point, on the
        GPU (triangulation_processor.py:160-234).  Returns a new (4, m) array; the input is left
        untouched (np.copy at :208)."""
        if not damping_factor:
            damping_factor = self.damping_factor
        if not iteration:
            iteration = self.iteration
        num_pts = matched_pairs[0].shape[1]
        num_views = len(projs)
        refined = np.copy(init_3d_pts)
        if num_pts == 0:
            return refined
        proj_arr = np.stack([np.asarray(p, dtype=np.float64) for p in projs]).reshape(num_views, 12)
        # point-major CSR: every point is seen by all n views, in view order
        rowptr = (np.arange(num_pts + 1, dtype=np.int64) * num_views).astype(np.int32)
        obs_view = np.tile(np.arange(num_views, dtype=np.int32), num_pts)
        obs_uv = np.stack([np.asarray(mp, dtype=np.float64)[0:2, :].T for mp in matched_pairs], axis=1)
        pts = np.ascontiguousarray(np.asarray(refined, dtype=np.float64)[0:3, :].T)
        out = self.engine.triangulate_refine(proj_arr, rowptr, obs_view, obs_uv.reshape(-1, 2), pts,
                                             damping_factor, iteration)
        refined[0:3, :] = out.T
        return refined

    # ------------------------------------------------------------------------------------------
    def construct_jacobian_matrix(self, tri_3d_pt, projs, num_views):
        """(2 n) x 3 Jacobian of the projections of one point (triangulation_processor.py:237-271).
        Single-observation host helper kept for API compatibility; the batched kernels evaluate the
        same expression on the device."""
        x = np.asarray(tri_3d_pt, dtype=float).reshape(-1)
        jac = np.zeros((num_views * 2, 3))
        for v in range(num_views):
            proj = np.asarray(projs[v], dtype=float)
            p = proj @ x
            jac[2 * v] = (p[2] * proj[0, 0:3] - p[0] * proj[2, 0:3]) / p[2] ** 2
            jac[2 * v + 1] = (p[2] * proj[1, 0:3] - p[1] * proj[2, 0:3]) / p[2] ** 2
        return jac
