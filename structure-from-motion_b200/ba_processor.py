"""BaProcessor with the reference's interface (ba_processor.py:21-440).

`_BaProcessor__execute_bundle_adjustment` -- the reference's dense NumPy Schur-complement
Gauss-Newton (ba_processor.py:274-440) -- is replaced by one call into libsfmba: the visibility
index is built bit-exactly from the track tables, the problem is packed into SoA arrays, the GPU
runs `iteration` fixed-lambda steps, and the results are written back through
`View.update_cam_pose` and in place into `tri_pts` exactly as the reference does.

The sequential host state machine `process` (one image / epipolar / camera-pose states, ba_processor.py:43-270) is NOT
restated here: it is control plane that stays the reference's.  The class is named like the reference's on purpose --
`self.__execute_bundle_adjustment()` inside the reference's `process` is name-mangled to
`_BaProcessor__execute_bundle_adjustment`, so

    import ba_processor as ref                       # the reference module, unchanged
    class BaProcessor(sfm_b200.BaProcessor, ref.BaProcessor):
        pass

runs the reference's `process` with the bundle adjustment below.
"""
from __future__ import annotations

import math

import numpy as np

from . import _capi
from .engine import get_engine
from .index import problem_from_tracks
from .utils import quaternion_to_rotation_unchecked, verify_rotation_mat

BA_DEBUG = False


class BaProcessor:
    def __init__(self, view_processor, key_tracker, epi_processor, tri_processor, campose_processor,
                 filter_size=10, iteration=3, damping_factor=5, engine=None, verbose=True):
        self.curr_data_idx = 0
        self.filter_size = filter_size          # ba_processor.py:33-40
        self.iteration = iteration
        self.damping_factor = damping_factor
        self.view_processor = view_processor
        self.key_tracker = key_tracker
        self.epi_processor = epi_processor
        self.tri_processor = tri_processor
        self.campose_processor = campose_processor
        self._engine = engine
        self.verbose = verbose                   # the reference always prints (ba_processor.py:418)
        self.last_result = None                  # SolveResult of the most recent BA call
        # engine knobs (reference behaviour by default)
        self.dvdc_sign_quirk = True
        self.pcg_tol = 1e-8
        self.use_gpu_index = True                # visibility CSR built on the device (bit-exact with is_visible)

    @property
    def engine(self):
        if self._engine is None:
            self._engine = get_engine()
        return self._engine

    # ------------------------------------------------------------------------------------------
    def __execute_bundle_adjustment(self):
        """Global refinement of all poses and points on the GPU (ba_processor.py:274-440)."""
        views = self.view_processor.view_list
        tri_pts = self.tri_processor.tri_pts
        view_num = len(views)
        tri_num = tri_pts.shape[1]
        prob = problem_from_tracks(views, tri_pts, self.key_tracker, self.engine, self.use_gpu_index)
        init_cams = prob.cams.copy()
        init_pts = prob.pts.copy()
        self.last_result = self.engine.ba_solve_host(prob, lam=self.damping_factor, iterations=self.iteration,
                                                     pcg_tol=self.pcg_tol, dvdc_sign_quirk=self.dvdc_sign_quirk)
        if self.last_result.flags & _capi.FLAG_PCG_MAXITER:
            # The reference solves the reduced system exactly (ba_processor.py:382).  A PCG solve that stopped on its
            # iteration cap is not that: redo the call from the same start with the direct (dense Cholesky) solver -- at
            # most 7 * filter_size unknowns here.
            prob.cams[...] = init_cams
            prob.pts[...] = init_pts
            self.last_result = self.engine.ba_solve_host(prob, lam=self.damping_factor, iterations=self.iteration,
                                                         dvdc_sign_quirk=self.dvdc_sign_quirk,
                                                         solver=_capi.SOLVER_DENSE_CHOLESKY)
        # Write back.  The quaternion is renormalised in f64 on the host before R is rebuilt so the
        # result passes the reference's one-sided 1e-8 validity check (utils.py:101-105).
        for i in range(view_num):
            q = prob.cams[i, 3:7]
            q = q / math.sqrt(float(np.sum(np.square(q))))
            rot = quaternion_to_rotation_unchecked(q)
            if not verify_rotation_mat(rot):
                raise ValueError('{} : Invalid output rotation matrix \n {}'.format('convert_quaternion_to_rotation', rot))
            views[i].update_cam_pose(rot, prob.cams[i, 0:3].reshape(3, 1).copy())
        tri_pts[0:3, :] = prob.pts.T               # in place: other code holds this array (ba_processor.py:416)
        if self.verbose:
            for i in range(view_num):
                d = math.sqrt(float(np.sum(np.square(init_cams[i, 0:3] - prob.cams[i, 0:3]))))
                print('DEBUG: {}-th view loc distance changes {} unit'.format(i, d))
            moved = np.sqrt(np.sum(np.square(init_pts - prob.pts), axis=1))
            for j in np.flatnonzero(moved >= 5):
                print('DEBUG: {}-th pt loc changes more than 5 unit, {} unit'.format(j, moved[j]))
        _ = tri_num
