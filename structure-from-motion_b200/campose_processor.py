"""CamposeProcessor with the reference's interface (campose_processor.py:11-808).

The nonlinear PnP refinement (campose_processor.py:308-459) runs on the GPU through
sfmba_pnp_refine_batched, including the reference's row-packing and dv/dC quirks.  The RANSAC-DLT
initialiser (campose_processor.py:485-633) evaluates all its hypotheses on the GPU (sfmba_pnp_ransac,
SURVEY.md section 8(f) row 3) from sample indices drawn with Python's `random.sample` exactly like
the reference.  Essential-matrix decomposition and the cheirality vote are a handful of 3x3
operations and stay host code.
"""
from __future__ import annotations

import logging
import math
import random

import numpy as np

from .engine import get_engine
from .utils import convert_rotation_to_quaternion


class CamposeProcessor:
    def __init__(self, ransac_config, damping_factor, iteration, engine=None):
        self.ransac_config = ransac_config          # campose_processor.py:12-26
        self.damping_factor = damping_factor
        self.iteration = iteration
        self._engine = engine

    @property
    def engine(self):
        if self._engine is None:
            self._engine = get_engine()
        return self._engine

    # ------------------------------------------------------------------------------------------
    def extract_cam_pose_from_essential_mat(self, esse_mat):
        """Four (R, C) candidates from E (campose_processor.py:29-99)."""
        w = np.array([[0.0, -1.0, 0.0], [1.0, 0.0, 0.0], [0.0, 0.0, 1.0]])
        u, _, vh = np.linalg.svd(esse_mat)
        c1 = u[:, 2].reshape(-1, 1)
        c2 = -c1
        r1 = u @ w @ vh
        r2 = u @ w.T @ vh
        if np.linalg.det(r1) < 0:
            r1 = -r1
        if np.linalg.det(r2) < 0:
            r2 = -r2
        return r1.transpose(), r2.transpose(), c1, c2

    # ------------------------------------------------------------------------------------------
    def disambiguate_cam_pose_four(self, ref_proj, projs_four, tri_3d_pts_four):
        """Pick the candidate with most points in front of both cameras (campose_processor.py:102-136)."""
        most_valid_count = 0
        best_idx = 0
        most_valid_indices = []
        for idx in range(4):
            valid = self.evalulate_cam_pose_cheirality(ref_proj, projs_four[idx], tri_3d_pts_four[idx])
            if len(valid) > most_valid_count:
                most_valid_count = len(valid)
                best_idx = idx
                most_valid_indices = valid
        return best_idx, most_valid_indices

    # ------------------------------------------------------------------------------------------
    def evalulate_cam_pose_cheirality(self, proj_1, proj_2, tri_3d_pts):
        """Indices of the points with positive depth in both cameras (campose_processor.py:139-189)."""
        pts = np.asarray(tri_3d_pts, dtype=float)
        z1 = (np.asarray(proj_1, dtype=float) @ pts)[2]
        z2 = (np.asarray(proj_2, dtype=float) @ pts)[2]
        return [int(i) for i in np.flatnonzero((z1 > 0) & (z2 > 0))]

    # ------------------------------------------------------------------------------------------
    def estimate_cam_pose_pnp(self, key_2d_pts, tri_3d_pts, intrinsic_mat, ransac_config=None, damping_factor=None,
                              iteration=None):
        """RANSAC-DLT then nonlinear refinement on the inliers (campose_processor.py:192-246)."""
        if not ransac_config:
            ransac_config = self.ransac_config
        if not damping_factor:
            damping_factor = self.damping_factor
        if not iteration:
            iteration = self.iteration
        inlier_indices, ini_rot, ini_loc = self.linear_estimate_cam_pose_pnp(key_2d_pts, tri_3d_pts, intrinsic_mat,
                                                                              ransac_config)
        ref_rot, ref_loc = self.nonlinear_estimate_cam_pose_pnp(key_2d_pts[:, inlier_indices],
                                                                tri_3d_pts[:, inlier_indices], intrinsic_mat, ini_rot,
                                                                ini_loc, damping_factor, iteration)
        return inlier_indices, ref_rot, ref_loc

    # ------------------------------------------------------------------------------------------
    def linear_estimate_cam_pose_pnp(self, key_2d_pts, tri_3d_pts, intrinsic_mat, ransac_config=None,
                                     lapack_sign_quirk=True):
        """RANSAC over 6-point DLT poses (campose_processor.py:249-305, 485-633).

        The six indices of every hypothesis are drawn here with random.sample, one draw per iteration in the
        reference's order (campose_processor.py:531), so Python's global random stream is consumed exactly as
        the reference consumes it.  All hypotheses are then evaluated at once on the GPU (sfmba_pnp_ransac:
        12x12 null vector, rotation re-orthogonalisation, inlier count over every point, first strict maximum).

        lapack_sign_quirk: campose_processor.py:620-631 returns (R, C) or (R, -C) depending on the sign LAPACK
        happens to give the last right singular vector (about half of the hypotheses each).  True (default)
        reproduces that outcome hypothesis by hypothesis -- the sign, and only the sign, is read off a batched
        numpy.linalg.svd of the same 12x12 systems on the host; False uses the geometrically consistent sign
        everywhere (pure device path, usually a better winner than the reference finds)."""
        if not ransac_config:
            ransac_config = self.ransac_config
        if key_2d_pts.shape[1] != tri_3d_pts.shape[1]:
            logging.warning('%s : different numbers of key points and of triangulated points',
                            self.__class__.__name__)
            raise ValueError('key pts num - triangulated pts num : {} - {}'.format(key_2d_pts.shape[1],
                                                                                   tri_3d_pts.shape[1]))
        num_pts = key_2d_pts.shape[1]
        if num_pts < 6:
            logging.warning('%s : required equal or more than six points', self.__class__.__name__)
            raise ValueError('required equal or more than six points {}'.format(num_pts))
        key = np.asarray(key_2d_pts, dtype=float)
        tri = np.asarray(tri_3d_pts, dtype=float)
        k = np.asarray(intrinsic_mat, dtype=float)
        samples = np.array([random.sample(range(num_pts), 6) for _ in range(ransac_config.iteration)],
                           dtype=np.int32).reshape(-1, 6)
        sign = self._lapack_loc_signs(key, tri, k, samples) if lapack_sign_quirk else None
        res = self.engine.pnp_ransac(key.T, tri.T, k, samples, ransac_config.inlier_threshold, sign)
        inliers = [int(i) for i in np.flatnonzero(res['mask'])]
        return inliers, res['rot'], res['loc']

    @staticmethod
    def _lapack_loc_signs(key, tri, k, samples):
        """+1 / -1 per hypothesis: the sign of det(rot_t_noise) for the last right singular vector as the host's
        LAPACK returns it (campose_processor.py:616-631).  One batched SVD; nothing else of the model is used."""
        if samples.shape[0] == 0:
            return np.zeros(0, dtype=np.int32)
        p2 = np.linalg.inv(k) @ key                                   # (3, n)
        a = p2[:, samples]                                            # (3, H, 6)
        x = tri[0:3][:, samples]                                      # (3, H, 6)
        H = samples.shape[0]
        w = np.zeros((H, 12, 12))
        zx = (a[2][None] * x).transpose(1, 2, 0)                      # (H, 6, 3)
        w[:, 0::2, 0:3] = zx
        w[:, 0::2, 3] = a[2]
        w[:, 0::2, 8:11] = -(a[0][None] * x).transpose(1, 2, 0)
        w[:, 0::2, 11] = -a[0]
        w[:, 1::2, 4:7] = zx
        w[:, 1::2, 7] = a[2]
        w[:, 1::2, 8:11] = -(a[1][None] * x).transpose(1, 2, 0)
        w[:, 1::2, 11] = -a[1]
        vh = np.linalg.svd(w)[2]
        m = vh[:, -1, :].reshape(H, 3, 4)[:, :, 0:3]
        return np.where(np.linalg.det(m) < 0, -1, 1).astype(np.int32)

    # ------------------------------------------------------------------------------------------
    def nonlinear_estimate_cam_pose_pnp(self, key_2d_pts, tri_3d_pts, intrinsic_mat, init_rot, init_loc,
                                        damping_factor=None, iteration=None, row_packing_quirk=True,
                                        dvdc_sign_quirk=True):
        """Damped Gauss-Newton on [C, q] with fixed lambda and iteration count, on the GPU
        (campose_processor.py:308-459).  Returns (rot (3,3), loc (3,1)).

        row_packing_quirk / dvdc_sign_quirk default to the reference's behaviour
        (campose_processor.py:404-405 and 802-804)."""
        if not damping_factor:
            damping_factor = self.damping_factor
        if not iteration:
            iteration = self.iteration
        if key_2d_pts.shape[1] != tri_3d_pts.shape[1]:
            logging.warning('%s : different numbers of key points and of triangulated points',
                            self.__class__.__name__)
            raise ValueError('key pts num - triangulated pts num : {} - {}'.format(key_2d_pts.shape[1],
                                                                                   tri_3d_pts.shape[1]))
        init_rot = np.asarray(init_rot, dtype=float)
        convert_rotation_to_quaternion(init_rot)       # campose_processor.py:361: ValueError if invalid
        key = np.asarray(key_2d_pts, dtype=float)
        tri = np.asarray(tri_3d_pts, dtype=float)
        num_pts = key.shape[1]
        if num_pts and not np.all(tri[3] == 1.0):
            raise ValueError('tri_3d_pts must be homogeneous with last row 1 (TriPt)')
        cam = np.linalg.inv(np.asarray(intrinsic_mat, dtype=float)) @ key     # campose_processor.py:393-394
        obs_xy = (cam[0:2] / cam[2:3]).T
        rowptr = np.array([0, num_pts], dtype=np.int32)
        rot, loc = self.engine.pnp_refine(rowptr, obs_xy, tri[0:3].T, init_rot.reshape(1, 3, 3),
                                          np.asarray(init_loc, dtype=float).reshape(1, 3), damping_factor, iteration,
                                          row_packing_quirk, dvdc_sign_quirk)
        return rot[0], loc[0].reshape(3, 1)

    # ------------------------------------------------------------------------------------------
    def construct_jacobian_matrix(self, rot, loc, pt_3d, dvdc_sign_quirk=True):
        """2 x 7 Jacobian of the normalised projection w.r.t. [C, q] for one point
        (campose_processor.py:462-482), evaluated by the device kernel (K1)."""
        from .engine import BAProblem
        rot = np.asarray(rot, dtype=float)
        q = convert_rotation_to_quaternion(rot).reshape(4)     # campose_processor.py:464
        x = np.asarray(pt_3d, dtype=float).reshape(-1)[0:3]
        prob = BAProblem(cams=np.concatenate([np.asarray(loc, dtype=float).reshape(3), q])[None, :],
                         pts=x[None, :], obs_uv=np.zeros((2, 1), dtype=np.float32), obs_cam=np.zeros(1, dtype=np.int32),
                         obs_pt=np.zeros(1, dtype=np.int32), rowptr=np.array([0, 1], dtype=np.int32))
        _, jp, _, _ = self.engine.residual_jacobian_f64(prob, dvdc_sign_quirk)
        return jp[0]
