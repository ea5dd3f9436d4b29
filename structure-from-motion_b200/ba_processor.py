"""BaProcessor with the reference's interface (ba_processor.py:21-440).

`_BaProcessor__execute_bundle_adjustment` -- the reference's dense NumPy Schur-complement
Gauss-Newton (ba_processor.py:274-440) -- is replaced by one call into libsfmba: the visibility
index is built bit-exactly from the track tables, the problem is packed into SoA arrays, the GPU
runs `iteration` fixed-lambda steps, and the results are written back through
`View.update_cam_pose` and in place into `tri_pts` exactly as the reference does.

`process` is the sequential host state machine (one image / epipolar / camera-pose states); it only
orchestrates the collaborating processors and stays on the host.
"""
from __future__ import annotations

import math
import sys

import numpy as np

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
        self.use_gpu_index = False

    @property
    def engine(self):
        if self._engine is None:
            self._engine = get_engine()
        return self._engine

    # ------------------------------------------------------------------------------------------
    def process(self, img, k):
        """Feed one image (ba_processor.py:43-270)."""
        if self.curr_data_idx >= self.filter_size:
            print('Bundle Adjustment processor is full')
            return
        vp, kt, tp, cp = self.view_processor, self.key_tracker, self.tri_processor, self.campose_processor
        view = vp.generate_view(img, self.curr_data_idx, k)
        kt.add_new_view(view, vp.view_list)
        vp.add_view(view)
        views = vp.view_list
        idx = self.curr_data_idx
        if idx == 0:
            print('In one image state')
            views[0].is_valid = True
        elif idx == 1:
            print('In epipolar state')
            self._epipolar_state(views, k)
        else:
            print('In cam pose state')
            self._campose_state(views, idx)
            self.__execute_bundle_adjustment()
        self.curr_data_idx += 1

    def _epipolar_state(self, views, k):
        """Two-view initialisation (ba_processor.py:63-135)."""
        kt, tp, cp, ep = self.key_tracker, self.tri_processor, self.campose_processor, self.epi_processor
        pairs, r_idx, q_idx = kt.generate_matched_pairs(0, 1, views)
        ep.determine_fundamental_mat(pairs)
        ep.extract_essential_mat(views[0].k, views[1].k)
        r1, r2, c1, c2 = cp.extract_cam_pose_from_essential_mat(ep.esse_mat)
        rots = [r1, r1, r2, r2]
        locs = [c1, c2, c1, c2]
        cand_projs = [k @ np.hstack((r.T, -r.T @ c)) for r, c in zip(rots, locs)]
        ref_proj = views[0].cam_proj
        cand_pts = [tp.linear_triangulate([ref_proj, p], pairs) for p in cand_projs]
        best, valid = cp.disambiguate_cam_pose_four(ref_proj, cand_projs, cand_pts)
        views[1].update_cam_pose(rots[best], locs[best])
        valid = np.array(valid)[np.newaxis, :]
        sel_pairs = [np.take_along_axis(pairs[0], valid, axis=1), np.take_along_axis(pairs[1], valid, axis=1)]
        sel_pts = np.take_along_axis(cand_pts[best], valid, axis=1)
        refined = tp.nonlinear_triangulate(sel_pts, [ref_proj, cand_projs[best]], sel_pairs)
        tri_indices = np.arange(0, refined.shape[1], dtype=int)[np.newaxis, :]
        kt.track_list[0].update_usage(np.take(r_idx, valid), tri_indices)
        kt.track_list[1].update_usage(np.take(q_idx, valid), tri_indices)
        views[1].is_valid = True
        views[1].ref_idx = 0
        tp.add_tri_pt(refined)

    def _campose_state(self, views, cur_idx):
        """Register a further view by PnP and triangulate its new points (ba_processor.py:137-264)."""
        kt, tp, cp = self.key_tracker, self.tri_processor, self.campose_processor
        cur = views[cur_idx]
        best_idx = kt.find_best_view(cur_idx)
        best = views[best_idx]
        cur.ref_idx = best_idx
        pairs, b_key_idx, _ = kt.generate_matched_pairs(best_idx, cur_idx, views)
        b_tri_idx, b_tri_val = kt.track_list[best_idx].extract_constructed_points()
        if b_tri_idx.shape[1] != tp.tri_pts.shape[1]:
            sys.exit('ERROR: even best_view_idx assumption does not work !!!')
        used_val, b_key_pos, b_tri_pos = np.intersect1d(b_key_idx, b_tri_idx, return_indices=True)
        used_val = used_val[np.newaxis, :]
        b_key_pos = b_key_pos[np.newaxis, :]
        tri_sel = np.take_along_axis(b_tri_val, b_tri_pos[np.newaxis, :], axis=1)
        known_pts = np.take_along_axis(tp.tri_pts, tri_sel, axis=1)
        cur_keys = np.take_along_axis(pairs[1], b_key_pos, axis=1)
        _, c_rot, c_loc = cp.estimate_cam_pose_pnp(cur_keys, known_pts, cur.k)
        cur.update_cam_pose(c_rot, c_loc)
        cur.is_valid = True
        # new points: matches whose key in the best view has not been triangulated yet
        unused = kt.track_list[best_idx].extract_unconstructed_points()
        new_val, new_pos, _ = np.intersect1d(b_key_idx, unused, return_indices=True)
        new_val = new_val[np.newaxis, :]
        both, _, _ = np.intersect1d(new_val, used_val, return_indices=True)
        if np.any(both) != False:  # noqa: E712  (same test as ba_processor.py:235)
            sys.exit('ERROR: the intersect of unused and used inters sets is NOT empty')
        new_pos = new_pos[np.newaxis, :]
        new_pairs = [np.take_along_axis(pairs[0], new_pos, axis=1), np.take_along_axis(pairs[1], new_pos, axis=1)]
        new_pts = tp.triangulate([best.cam_proj, cur.cam_proj], new_pairs)
        n0 = tp.tri_pts.shape[1]
        tri_indices = np.arange(n0, n0 + new_pts.shape[1], dtype=int)[np.newaxis, :]
        best_track = kt.track_list[best_idx]
        best_track.update_usage(new_val, tri_indices)
        cur_keys_idx = np.take(best_track.table[cur_idx, :], new_val)
        kt.track_list[cur_idx].update_usage(cur_keys_idx, tri_indices)
        tp.add_tri_pt(new_pts)

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
