"""Shared helpers for the parity tests."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name))


def ba_case_from_npz(z, prefix=''):
    """Rebuild the inputs of one recorded BaProcessor.__execute_bundle_adjustment call."""
    g = lambda k: z[prefix + k]  # noqa: E731
    V = g('rot').shape[0]
    N = g('tri_pts').shape[1]
    diag = [g('diag_%d' % i) for i in range(V)]
    keys = [g('keys_%d' % i) for i in range(V)]
    return dict(V=V, N=N, diag=diag, keys=keys, rot=g('rot'), loc=g('loc'), k=g('k'), tri_pts=g('tri_pts'),
                vis=g('vis'), jp0=g('jp0'), jx0=g('jx0'), b0=g('b0'), f0=g('f0'), out_rot=g('out_rot'),
                out_loc=g('out_loc'), out_tri_pts=g('out_tri_pts'), iteration=int(g('iteration')),
                damping_factor=float(g('damping_factor')))


class FakeKey:
    def __init__(self, pt):
        self.pt = (float(pt[0]), float(pt[1]))


class FakeView:
    """Stands in for view_processor.View: the fields BA reads and update_cam_pose (view_processor.py:61-69)."""

    def __init__(self, rot, loc, k, keys):
        self.rot = np.array(rot, dtype=float)
        self.loc = np.array(loc, dtype=float).reshape(3, 1)
        self.k = np.array(k, dtype=float)
        self.key_pts = [FakeKey(p) for p in keys]
        self.update_cam_pose(self.rot, self.loc)

    def update_cam_pose(self, rot, loc):
        self.rot = rot
        self.loc = loc
        self.cam_pose = np.hstack((self.rot, self.loc))
        self.cam_proj = self.k @ np.hstack((self.rot.T, self.rot.T @ -self.loc))


class FakeTrack:
    def __init__(self, table):
        self.table = table


class FakeTracker:
    def __init__(self, diag_rows):
        V = len(diag_rows)
        self.track_list = []
        for i, row in enumerate(diag_rows):
            t = np.full((V, row.shape[0]), -1, dtype=int)
            t[i, :] = row
            self.track_list.append(FakeTrack(t))


class FakeViewProcessor:
    def __init__(self, views):
        self.view_list = views


def problem_from_case(case):
    """BAProblem (product packing) of a recorded case."""
    import sfm_b200
    views = [FakeView(case['rot'][i], case['loc'][i], case['k'][i], case['keys'][i]) for i in range(case['V'])]
    tracker = FakeTracker(case['diag'])
    return sfm_b200.problem_from_tracks(views, case['tri_pts'], tracker), views, tracker


def close(a, b, tol):
    """|a - b| <= tol * max(1, |b|) elementwise (SURVEY.md section 7 'hard parts')."""
    a = np.asarray(a, dtype=float)
    b = np.asarray(b, dtype=float)
    return bool(np.all(np.abs(a - b) <= tol * np.maximum(1.0, np.abs(b))))


def max_rel(a, b):
    a = np.asarray(a, dtype=float)
    b = np.asarray(b, dtype=float)
    return float(np.max(np.abs(a - b) / np.maximum(1.0, np.abs(b)))) if a.size else 0.0
