"""BAL ("Bundle Adjustment in the Large") text files <-> BAProblem  (SURVEY.md section 8(f) row 4).

The reference has no file format for bundle-adjustment problems; this is the wire format that lets the
engine run real BAL problems (Ladybug, Venice, Final ...) instead of BAL-shaped synthetics.

File layout (whitespace separated):
    <num_cameras> <num_points> <num_observations>
    <camera_index> <point_index> <x> <y>            x num_observations
    <camera parameter>                               x 9 per camera: Rodrigues r(3), t(3), f, k1, k2
    <point coordinate>                               x 3 per point
BAL camera model:  P = R(r) X + t;  p = -P_xy / P_z;  pixel = f (1 + k1 |p|^2 + k2 |p|^4) p.

Conversion to the reference's camera model (7 parameters: centre C + quaternion (w,x,y,z), known intrinsics,
residual in normalised camera coordinates, ba_processor.py:285-288, 325-349):
  * BAL cameras look down -z.  With S = diag(-1, -1, 1) (a half turn about z) the point in the reference's
    camera frame is p_c = S P, so that p_c,xy / p_c,z = -P_xy / P_z.  Turning the world by the same S
    (X' = S X) keeps rotations close to the identity close to the identity:
        p_c = (S R S) X' + S t   =>   rot = (S R S)^T (camera-to-world),  C = -rot (S t)
    (p_c,z is negative for points in front of a BAL camera; the projection does not care.)
  * f, k1, k2 are held fixed (the reference's intrinsics are known): every observation is undistorted
    once on the host (Newton on the radius) and handed over as the undistorted pixel f p (float32, like
    cv2.KeyPoint.pt) with K^-1 = diag(1/f, 1/f, 1) per camera (ba_processor.py:339-342).
  * observations are sorted point-major / view-minor (ba_processor.py:304-306).
`problem_to_bal` inverts all of it, so a solve can be written back as a BAL file.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from .engine import BAProblem

_S = np.array([-1.0, -1.0, 1.0])


@dataclass
class BalFile:
    cams: np.ndarray        # (V, 9) r(3), t(3), f, k1, k2
    pts: np.ndarray         # (N, 3)
    obs_cam: np.ndarray     # (m,) int32, file order
    obs_pt: np.ndarray      # (m,) int32
    obs_xy: np.ndarray      # (m, 2) pixels

    @property
    def num_cams(self):
        return self.cams.shape[0]

    @property
    def num_pts(self):
        return self.pts.shape[0]

    @property
    def num_obs(self):
        return self.obs_cam.shape[0]


def read_bal(path) -> BalFile:
    """Parse a BAL text file (plain or .bz2 / .gz)."""
    p = str(path)
    if p.endswith('.bz2'):
        import bz2
        with bz2.open(p, 'rt') as f:
            data = np.array(f.read().split(), dtype=np.float64)
    elif p.endswith('.gz'):
        import gzip
        with gzip.open(p, 'rt') as f:
            data = np.array(f.read().split(), dtype=np.float64)
    else:
        data = np.fromfile(p, dtype=np.float64, sep=' ')
    if data.size < 3:
        raise ValueError('not a BAL file: %s' % p)
    V, N, m = int(data[0]), int(data[1]), int(data[2])
    want = 3 + 4 * m + 9 * V + 3 * N
    if V < 0 or N < 0 or m < 0 or data.size != want:
        raise ValueError('BAL file %s: header says %d cameras / %d points / %d observations = %d numbers, found %d'
                         % (p, V, N, m, want, data.size))
    obs = data[3:3 + 4 * m].reshape(m, 4)
    cam_idx = obs[:, 0].astype(np.int32)
    pt_idx = obs[:, 1].astype(np.int32)
    if m and (cam_idx.min() < 0 or cam_idx.max() >= V or pt_idx.min() < 0 or pt_idx.max() >= N):
        raise ValueError('BAL file %s: observation index out of range' % p)
    cams = data[3 + 4 * m:3 + 4 * m + 9 * V].reshape(V, 9).copy()
    pts = data[3 + 4 * m + 9 * V:].reshape(N, 3).copy()
    return BalFile(cams=cams, pts=pts, obs_cam=cam_idx, obs_pt=pt_idx, obs_xy=obs[:, 2:4].copy())


def write_bal(path, bal: BalFile) -> None:
    """Write a BAL text file (%.16e, so a read-back is bit-exact)."""
    with open(str(path), 'w') as f:
        f.write('%d %d %d\n' % (bal.num_cams, bal.num_pts, bal.num_obs))
        if bal.num_obs:
            lines = ['%d %d %.16e %.16e' % (c, p, x, y)
                     for c, p, (x, y) in zip(bal.obs_cam.tolist(), bal.obs_pt.tolist(), bal.obs_xy.tolist())]
            f.write('\n'.join(lines))
            f.write('\n')
        np.savetxt(f, bal.cams.reshape(-1, 1), fmt='%.16e')
        np.savetxt(f, bal.pts.reshape(-1, 1), fmt='%.16e')


# ---- rotations -----------------------------------------------------------------------------------------
def rodrigues_to_rot(r: np.ndarray) -> np.ndarray:
    """(n,3) angle-axis -> (n,3,3)."""
    r = np.asarray(r, dtype=np.float64).reshape(-1, 3)
    th = np.linalg.norm(r, axis=1)
    small = th < 1e-12
    k = r / np.where(small, 1.0, th)[:, None]
    K = np.zeros((r.shape[0], 3, 3))
    K[:, 0, 1], K[:, 0, 2] = -k[:, 2], k[:, 1]
    K[:, 1, 0], K[:, 1, 2] = k[:, 2], -k[:, 0]
    K[:, 2, 0], K[:, 2, 1] = -k[:, 1], k[:, 0]
    R = np.eye(3)[None] + np.sin(th)[:, None, None] * K + (1 - np.cos(th))[:, None, None] * (K @ K)
    if small.any():
        # first order: I + [r]x
        Ks = np.zeros((int(small.sum()), 3, 3))
        rs = r[small]
        Ks[:, 0, 1], Ks[:, 0, 2] = -rs[:, 2], rs[:, 1]
        Ks[:, 1, 0], Ks[:, 1, 2] = rs[:, 2], -rs[:, 0]
        Ks[:, 2, 0], Ks[:, 2, 1] = -rs[:, 1], rs[:, 0]
        R[small] = np.eye(3)[None] + Ks
    return R


def rot_to_quat(R: np.ndarray) -> np.ndarray:
    """(n,3,3) -> (n,4) unit quaternion (w,x,y,z), w >= 0, in the convention of utils.py:64-97 (the matrix
    that quaternion_to_rotation rebuilds is R).  All four branches, so half turns are fine -- the reference's
    own conversion (utils.py:28-60) only has the qw branch."""
    R = np.asarray(R, dtype=np.float64).reshape(-1, 3, 3)
    n = R.shape[0]
    q = np.empty((n, 4))
    t = np.stack([1 + R[:, 0, 0] + R[:, 1, 1] + R[:, 2, 2], 1 + R[:, 0, 0] - R[:, 1, 1] - R[:, 2, 2],
                  1 - R[:, 0, 0] + R[:, 1, 1] - R[:, 2, 2], 1 - R[:, 0, 0] - R[:, 1, 1] + R[:, 2, 2]], axis=1)
    br = np.argmax(t, axis=1)
    s = 2.0 * np.sqrt(np.maximum(t[np.arange(n), br], 1e-300))      # 4 |q_br|
    a = R[:, 2, 1] - R[:, 1, 2]      # 4 w x
    b = R[:, 0, 2] - R[:, 2, 0]      # 4 w y
    c = R[:, 1, 0] - R[:, 0, 1]      # 4 w z
    d = R[:, 0, 1] + R[:, 1, 0]      # 4 x y
    e = R[:, 0, 2] + R[:, 2, 0]      # 4 x z
    f = R[:, 1, 2] + R[:, 2, 1]      # 4 y z
    for k, cols in enumerate(((s / 4, a / s, b / s, c / s), (a / s, s / 4, d / s, e / s),
                              (b / s, d / s, s / 4, f / s), (c / s, e / s, f / s, s / 4))):
        sel = br == k
        for j in range(4):
            q[sel, j] = cols[j][sel]
    q *= np.where(q[:, 0:1] < 0, -1.0, 1.0)
    return q / np.linalg.norm(q, axis=1, keepdims=True)


def rot_to_rodrigues(R: np.ndarray) -> np.ndarray:
    q = rot_to_quat(R)
    sn = np.linalg.norm(q[:, 1:4], axis=1)
    ang = 2.0 * np.arctan2(sn, q[:, 0])
    scale = np.where(sn < 1e-300, 2.0, ang / np.where(sn < 1e-300, 1.0, sn))
    return q[:, 1:4] * scale[:, None]


def _quat_to_rot(q):
    from .synthetic import quat_to_rot
    return quat_to_rot(np.asarray(q, dtype=np.float64).reshape(-1, 4))


# ---- distortion ----------------------------------------------------------------------------------------
def undistort(pd: np.ndarray, k1: np.ndarray, k2: np.ndarray) -> np.ndarray:
    """Solve p (1 + k1 |p|^2 + k2 |p|^4) = pd for p, per row (Newton on the radius, 20 steps)."""
    rd = np.linalg.norm(pd, axis=1)
    r = rd.copy()
    for _ in range(20):
        r2 = r * r
        g = r * (1 + k1 * r2 + k2 * r2 * r2) - rd
        dg = 1 + 3 * k1 * r2 + 5 * k2 * r2 * r2
        r = r - g / np.where(dg == 0, 1.0, dg)
    scale = np.where(rd > 0, r / np.where(rd > 0, rd, 1.0), 1.0)
    return pd * scale[:, None]


def distort(p: np.ndarray, k1: np.ndarray, k2: np.ndarray) -> np.ndarray:
    r2 = np.sum(p * p, axis=1)
    return p * (1 + k1 * r2 + k2 * r2 * r2)[:, None]


# ---- conversion ----------------------------------------------------------------------------------------
def bal_to_problem(bal: BalFile):
    """BalFile -> (BAProblem in the reference's camera model, order) where `order` is the permutation that
    sorted the file's observations point-major / view-minor (problem observation i = file observation
    order[i])."""
    V = bal.num_cams
    order = np.lexsort((bal.obs_cam, bal.obs_pt)).astype(np.int64)
    obs_cam = bal.obs_cam[order].astype(np.int32)
    obs_pt = bal.obs_pt[order].astype(np.int32)
    rowptr = np.zeros(bal.num_pts + 1, dtype=np.int32)
    np.cumsum(np.bincount(obs_pt, minlength=bal.num_pts), out=rowptr[1:])
    Rb = rodrigues_to_rot(bal.cams[:, 0:3])
    Rp = _S[None, :, None] * Rb * _S[None, None, :]                     # S R S
    rot = np.transpose(Rp, (0, 2, 1))                                   # camera-to-world
    tp = bal.cams[:, 3:6] * _S[None, :]
    C = -np.einsum('nij,nj->ni', rot, tp)
    cams = np.concatenate([C, rot_to_quat(rot)], axis=1)
    pts = bal.pts * _S[None, :]
    f, k1, k2 = bal.cams[:, 6], bal.cams[:, 7], bal.cams[:, 8]
    oc = obs_cam
    pd = bal.obs_xy[order] / f[oc][:, None]
    pu = undistort(pd, k1[oc], k2[oc])
    obs_uv = (pu * f[oc][:, None]).T.astype(np.float32)                 # undistorted pixels, float32
    kinv = np.zeros((V, 9))
    kinv[:, 0] = 1.0 / f
    kinv[:, 4] = 1.0 / f
    kinv[:, 8] = 1.0
    return BAProblem(cams=cams, pts=pts, obs_uv=obs_uv, obs_cam=obs_cam, obs_pt=obs_pt, rowptr=rowptr, kinv=kinv), order


def problem_to_bal(prob: BAProblem, template: BalFile) -> BalFile:
    """Cameras and points of `prob` (after a solve) back into BAL parameters; intrinsics and observations are
    taken from `template`."""
    rot = _quat_to_rot(prob.cams[:, 3:7] / np.linalg.norm(prob.cams[:, 3:7], axis=1, keepdims=True))
    Rp = np.transpose(rot, (0, 2, 1))
    Rb = _S[None, :, None] * Rp * _S[None, None, :]
    tp = -np.einsum('nij,nj->ni', Rp, prob.cams[:, 0:3])
    cams = template.cams.copy()
    cams[:, 0:3] = rot_to_rodrigues(Rb)
    cams[:, 3:6] = tp * _S[None, :]
    return BalFile(cams=cams, pts=prob.pts * _S[None, :], obs_cam=template.obs_cam.copy(),
                   obs_pt=template.obs_pt.copy(), obs_xy=template.obs_xy.copy())


def bal_project(bal: BalFile) -> np.ndarray:
    """Pixel predictions (m,2) of the BAL camera model for every observation of the file, in file order."""
    R = rodrigues_to_rot(bal.cams[:, 0:3])
    P = np.einsum('nij,nj->ni', R[bal.obs_cam], bal.pts[bal.obs_pt]) + bal.cams[bal.obs_cam, 3:6]
    p = -P[:, 0:2] / P[:, 2:3]
    c = bal.cams[bal.obs_cam]
    return distort(p, c[:, 7], c[:, 8]) * c[:, 6:7]


def synthetic_bal(num_cams=8, num_pts=400, seed=0, noise_px=0.3) -> BalFile:
    """A small random BAL problem (cameras on an arc looking down -z at a point cloud), for tests."""
    rng = np.random.default_rng(seed)
    ang = np.linspace(-0.4, 0.4, num_cams)
    r = np.stack([rng.normal(0, 0.02, num_cams), ang, rng.normal(0, 0.02, num_cams)], axis=1)
    R = rodrigues_to_rot(r)
    centre = np.stack([3.0 * np.sin(ang), rng.normal(0, 0.1, num_cams), 3.0 * (1 - np.cos(ang))], axis=1)
    t = -np.einsum('nij,nj->ni', R, centre)
    f = rng.uniform(500, 900, num_cams)
    k1 = rng.normal(0, 0.05, num_cams)
    k2 = rng.normal(0, 0.01, num_cams)
    cams = np.concatenate([r, t, f[:, None], k1[:, None], k2[:, None]], axis=1)
    pts = np.stack([rng.uniform(-2, 2, num_pts), rng.uniform(-1.5, 1.5, num_pts), rng.uniform(-9, -5, num_pts)], axis=1)
    oc, op = [], []
    for j in range(num_pts):
        k = int(rng.integers(2, min(num_cams, 6) + 1))
        for c in sorted(rng.choice(num_cams, k, replace=False).tolist()):
            oc.append(c)
            op.append(j)
    perm = rng.permutation(len(oc))                                     # BAL files are not point-major
    bal = BalFile(cams=cams, pts=pts, obs_cam=np.array(oc, dtype=np.int32)[perm],
                  obs_pt=np.array(op, dtype=np.int32)[perm], obs_xy=np.zeros((len(oc), 2)))
    bal.obs_xy = bal_project(bal) + rng.normal(0, noise_px, (len(oc), 2))
    return bal
