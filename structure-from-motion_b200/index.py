"""Visibility / track index construction and problem packing (host side).

The reference asks `KeyTracker.is_visible(view_idx, tri_idx)` for every (point, view) pair in a
point-major, view-minor double loop (ba_processor.py:304-310); is_visible (key_tracker.py:198-204)
scans the diagonal row `track_list[v].table[v, :]` and applies np.any() to the *indices* it finds.
The index built here is bit-exact with that loop, including:
  * a point whose only matching key index in a view is 0 is NOT visible there,
  * with several matching keys the smallest key index is used.
`Engine.build_visibility` is the same construction on the GPU.
"""
from __future__ import annotations

import numpy as np

from .engine import BAProblem
from .utils import convert_rotation_to_quaternion


def build_visibility_host(diag_rows, num_pts: int):
    """diag_rows[v] = table[v, :] of view v (int, -1 = key not used).  Returns int32 arrays
    (obs_pt, obs_cam, obs_key, rowptr) in the reference's loop order."""
    V = len(diag_rows)
    t_all, v_all, k_all = [], [], []
    for v in range(V):
        row = np.asarray(diag_rows[v]).reshape(-1)
        k = np.flatnonzero((row >= 0) & (row < num_pts))
        t_all.append(row[k].astype(np.int64))
        v_all.append(np.full(k.shape[0], v, dtype=np.int64))
        k_all.append(k.astype(np.int64))
    if V == 0 or sum(a.shape[0] for a in t_all) == 0:
        z = np.zeros(0, dtype=np.int32)
        return z, z.copy(), z.copy(), np.zeros(num_pts + 1, dtype=np.int32)
    t = np.concatenate(t_all)
    v = np.concatenate(v_all)
    k = np.concatenate(k_all)
    order = np.lexsort((k, v, t))                      # point-major, view-minor, smallest key first
    t, v, k = t[order], v[order], k[order]
    group = t * V + v
    head = np.ones(group.shape[0], dtype=bool)
    head[1:] = group[1:] != group[:-1]
    several = np.zeros(group.shape[0], dtype=bool)
    several[:-1] = group[1:] == group[:-1]             # the head's group has more than one key
    keep = head & ((k > 0) | several)                  # np.any(indices) semantics
    obs_pt = t[keep].astype(np.int32)
    obs_cam = v[keep].astype(np.int32)
    obs_key = k[keep].astype(np.int32)
    rowptr = np.searchsorted(obs_pt, np.arange(num_pts + 1), side='left').astype(np.int32)
    return obs_pt, obs_cam, obs_key, rowptr


def key_xy(view) -> np.ndarray:
    """(n_keys, 2) float32 pixel coordinates of a view's key points (cv2.KeyPoint.pt is float32), built once per view
    and kept on the view object: every later bundle adjustment gathers from it instead of touching the key objects."""
    keys = view.key_pts
    cached = getattr(view, '_sfmba_key_xy', None)
    if cached is not None and cached[0] is keys and cached[1].shape[0] == len(keys):
        return cached[1]
    xy = None
    try:                                                   # cv2.KeyPoint lists convert in one C call
        import cv2
        if len(keys) and isinstance(keys[0], cv2.KeyPoint):
            xy = np.asarray(cv2.KeyPoint_convert(keys), dtype=np.float32).reshape(-1, 2)
    except Exception:
        xy = None
    if xy is None:
        xy = np.array([kp.pt for kp in keys], dtype=np.float32).reshape(-1, 2)
    try:
        view._sfmba_key_xy = (keys, xy)
    except Exception:
        pass
    return xy


def problem_from_tracks(views, tri_pts, key_tracker, engine=None, use_gpu_index: bool = True) -> BAProblem:
    """Pack the reference's objects into a BAProblem (ba_processor.py:285-294, 309-342).

    views[i] needs .rot (3,3), .loc (3,1), .k (3,3), .key_pts[k].pt; tri_pts is the (4,N) TriPt
    array; key_tracker.track_list[i].table is the per-view track table."""
    V = len(views)
    N = int(tri_pts.shape[1])
    diag = [np.asarray(key_tracker.track_list[i].table[i, :]) for i in range(V)]
    if use_gpu_index and engine is not None:
        obs_pt, obs_cam, obs_key, rowptr = engine.build_visibility(diag, N)
    else:
        obs_pt, obs_cam, obs_key, rowptr = build_visibility_host(diag, N)
    cams = np.zeros((V, 7))
    kinv = np.zeros((V, 9))
    for i, view in enumerate(views):
        # ba_processor.py:287-288: q from the view's rotation (validated, ValueError if invalid)
        cams[i, 0:3] = np.asarray(view.loc, dtype=float).reshape(3)
        cams[i, 3:7] = convert_rotation_to_quaternion(view.rot).reshape(4)
        kinv[i] = np.linalg.inv(np.asarray(view.k, dtype=float)).reshape(9)    # ba_processor.py:340
    m = obs_pt.shape[0]
    uv = np.zeros((2, m), dtype=np.float32)
    for i, view in enumerate(views):
        sel = np.flatnonzero(obs_cam == i)
        if sel.size:
            uv[:, sel] = key_xy(view)[obs_key[sel]].T          # one gather per view (ba_processor.py:336-338)
    pts = np.ascontiguousarray(np.asarray(tri_pts, dtype=float)[0:3, :].T)
    return BAProblem(cams=cams, pts=pts, obs_uv=uv, obs_cam=obs_cam, obs_pt=obs_pt, rowptr=rowptr, kinv=kinv,
                     obs_key=obs_key)
