"""Synthetic BAL-shaped bundle-adjustment problems and point-sharding.

"BAL-shaped" = BAL sizes and sparsity with the REFERENCE's camera model: 7 parameters per camera
(centre C + quaternion (w,x,y,z)), known intrinsics, residual in normalised camera coordinates
(ba_processor.py:285-288, 325-349).  Everything is seeded and deterministic.
"""
from __future__ import annotations

import numpy as np

from .engine import BAProblem

# name -> (cameras, points, observations, seed)        BASELINE.json configs / SURVEY.md section 8
SHAPES = {
    'ladybug': (49, 7776, 31843, 0),
    'venice': (1778, 993923, 5001946, 1),
    'final': (13682, 4456117, 28987644, 2),
    'tiny': (6, 300, 1100, 5),
    'small': (24, 3000, 13000, 4),
}


CAM_SPACING = 0.3      # distance between consecutive camera centres


def _quat_mul(a, b):
    w1, x1, y1, z1 = a[..., 0], a[..., 1], a[..., 2], a[..., 3]
    w2, x2, y2, z2 = b[..., 0], b[..., 1], b[..., 2], b[..., 3]
    return np.stack([w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                     w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2, w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2], axis=-1)


def _axis_angle_quat(axis, angle):
    axis = axis / np.linalg.norm(axis, axis=-1, keepdims=True)
    h = 0.5 * angle
    return np.concatenate([np.cos(h)[..., None], axis * np.sin(h)[..., None]], axis=-1)


def quat_to_rot(q):
    """(n,4) (w,x,y,z) -> (n,3,3), same polynomial as utils.py:82-91."""
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    r = np.empty((q.shape[0], 3, 3))
    r[:, 0, 0] = 1 - 2 * z * z - 2 * y * y
    r[:, 0, 1] = -2 * z * w + 2 * y * x
    r[:, 0, 2] = 2 * y * w + 2 * z * x
    r[:, 1, 0] = 2 * x * y + 2 * w * z
    r[:, 1, 1] = 1 - 2 * z * z - 2 * x * x
    r[:, 1, 2] = 2 * z * y - 2 * x * w
    r[:, 2, 0] = 2 * x * z - 2 * w * y
    r[:, 2, 1] = 2 * y * z + 2 * w * x
    r[:, 2, 2] = 1 - 2 * y * y - 2 * x * x
    return r


def _track_lengths(rng, N, m, V):
    """2 + geometric track lengths with total exactly m, each in [2, V]."""
    mean_extra = max(m / N - 2.0, 1e-3)
    p = 1.0 / (1.0 + mean_extra)
    L = 2 + rng.geometric(p, size=N) - 1
    L = np.minimum(L, V).astype(np.int64)
    diff = int(m - L.sum())
    while diff != 0:
        k = min(abs(diff), N)
        idx = rng.choice(N, size=k, replace=False)
        if diff > 0:
            ok = idx[L[idx] < V]
            L[ok] += 1
            diff -= ok.size
        else:
            ok = idx[L[idx] > 2]
            L[ok] -= 1
            diff += ok.size
    return L


def make_problem(V: int, N: int, m: int, seed: int = 0, noise_px: float = 0.5, focal: float = 569.0,
                 perturb: bool = True):
    """Returns (problem, truth) where problem is the perturbed BAProblem to optimise and truth a dict
    with the exact cameras / points."""
    if m < 2 * N or m > V * N:
        raise ValueError('need 2N <= m <= V*N')
    rng = np.random.default_rng(seed)
    i = np.arange(V, dtype=np.float64)
    C = np.stack([CAM_SPACING * i, 0.5 * np.sin(i / 50.0), 0.3 * np.cos(i / 70.0)], axis=1)
    yaw = np.deg2rad(25.0) * np.sin(i / 100.0)
    pitch = np.deg2rad(5.0) * np.sin(i / 37.0)
    q = _quat_mul(_axis_angle_quat(np.tile([0.0, 1.0, 0.0], (V, 1)), yaw),
                  _axis_angle_quat(np.tile([1.0, 0.0, 0.0], (V, 1)), pitch))
    R = quat_to_rot(q)                                          # camera-to-world

    L = _track_lengths(rng, N, m, V)
    rowptr = np.zeros(N + 1, dtype=np.int64)
    rowptr[1:] = np.cumsum(L)
    obs_cam = np.empty(m, dtype=np.int64)
    # every camera is the centre of ~N/V consecutive points -> every camera is observed
    centre = np.minimum((np.arange(N, dtype=np.int64) * V) // N, V - 1)
    for length in np.unique(L):
        idx = np.flatnonzero(L == length)
        W = int(min(V, max(2 * (int(np.ceil(0.75 * length)) + 2) + 1, length)))   # band width
        start = np.clip(centre[idx] - W // 2, 0, V - W)
        cpos = centre[idx] - start                              # centre camera inside the band
        off = np.arange(W)
        keys = rng.random((idx.size, W)) + 1.2 * np.abs(off[None, :] - cpos[:, None]) / W
        keys[np.arange(idx.size), cpos] = -1.0                  # always keep the centre camera
        take = int(length)
        sel = np.argpartition(keys, take - 1, axis=1)[:, :take] if take < W else np.tile(off, (idx.size, 1))
        cams = np.sort(start[:, None] + sel, axis=1)
        obs_cam[rowptr[idx][:, None] + np.arange(take)[None, :]] = cams
    obs_pt = np.repeat(np.arange(N, dtype=np.int64), L)

    # points in a slab in front of their centre camera; longer tracks sit farther away
    local = np.stack([rng.uniform(-4, 4, N), rng.uniform(-3, 3, N), rng.uniform(6, 16, N) + 0.3 * L], axis=1)
    X = C[centre] + np.einsum('nij,nj->ni', R[centre], local)

    d = X[obs_pt] - C[obs_cam]
    pcam = np.einsum('mkc,mk->mc', R[obs_cam], d)               # R^T d
    if not np.all(pcam[:, 2] > 0.5):
        raise AssertionError('synthetic point behind / too close to a camera')
    b = pcam[:, 0:2] / pcam[:, 2:3] + rng.normal(0.0, noise_px / focal, (m, 2))
    obs_uv = np.ascontiguousarray(b.T.astype(np.float32))

    truth = dict(cams=np.concatenate([C, q], axis=1), pts=X.copy())
    if perturb:
        Cn = C + rng.normal(0.0, 0.05, (V, 3))
        axis = rng.normal(0.0, 1.0, (V, 3))
        dq = _axis_angle_quat(axis, np.full(V, np.deg2rad(0.5)))
        qn = _quat_mul(dq, q)
        qn /= np.linalg.norm(qn, axis=1, keepdims=True)
        Xn = X + rng.normal(0.0, 1.0, (N, 3)) * (0.05 * local[:, 2:3])
    else:
        Cn, qn, Xn = C, q, X
    prob = BAProblem(cams=np.concatenate([Cn, qn], axis=1), pts=Xn, obs_uv=obs_uv, obs_cam=obs_cam.astype(np.int32),
                     obs_pt=obs_pt.astype(np.int32), rowptr=rowptr.astype(np.int32), kinv=None)
    return prob, truth


def make_shape(name: str, scale: float = 1.0, **kw):
    """Named BAL shape, optionally scaled down (cameras, points and observations by `scale`)."""
    V, N, m, seed = SHAPES[name]
    if scale != 1.0:
        V = max(3, int(round(V * scale)))
        N2 = max(8, int(round(N * scale)))
        m = max(2 * N2, min(int(round(m * scale)), V * N2))
        N = N2
    return make_problem(V, N, m, seed=kw.pop('seed', seed), **kw)


def shard_bounds(rowptr: np.ndarray, world: int):
    """Contiguous point ranges balanced by observation count (prefix sum over track lengths)."""
    N = rowptr.shape[0] - 1
    m = int(rowptr[-1])
    targets = (np.arange(world + 1, dtype=np.int64) * m) // world
    b = np.searchsorted(rowptr, targets, side='left').astype(np.int64)
    b[0], b[-1] = 0, N
    return np.maximum.accumulate(b)


def shard_problem(p: BAProblem, world: int, rank: int) -> BAProblem:
    """This rank's shard: cameras replicated, points [p0, p1) with their observations."""
    b = shard_bounds(p.rowptr, world)
    p0, p1 = int(b[rank]), int(b[rank + 1])
    o0, o1 = int(p.rowptr[p0]), int(p.rowptr[p1])
    return BAProblem(cams=p.cams.copy(), pts=p.pts[p0:p1].copy(), obs_uv=p.obs_uv[:, o0:o1].copy(),
                     obs_cam=p.obs_cam[o0:o1].copy(), obs_pt=p.obs_pt[o0:o1] - p0,
                     rowptr=p.rowptr[p0:p1 + 1] - o0, kinv=None if p.kinv is None else p.kinv.copy())


# ---- config C3 of BASELINE.json: batched nonlinear triangulation / nonlinear PnP (a18, a19) --------------------
def tri_problem(M: int, rng, P: int = 200):
    """M points with 2..6 observations each out of P projections on a slowly turning track.  Returns
    (projs (P,3,4), rowptr (M+1,) i32, obs_view (m,) i32, uv (m,2) pixels with 0.5 px noise, X0 (M,3) perturbed start)."""
    K = np.array([[569.0, 0, 640.0], [0, 569.0, 480.0], [0, 0, 1.0]])
    projs = np.zeros((P, 3, 4))
    for i in range(P):
        ang = 0.2 * np.sin(i / 30.0)
        R = np.array([[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]])
        C = np.array([[i * 0.2], [0.0], [0.0]])
        projs[i] = K @ np.hstack((R.T, -R.T @ C))
    n = rng.integers(2, 7, M)
    rowptr = np.zeros(M + 1, dtype=np.int32)
    rowptr[1:] = np.cumsum(n)
    first = rng.integers(0, P - 6, M)
    obs_view = (np.repeat(first, n) + (np.arange(rowptr[-1]) - np.repeat(rowptr[:-1], n))).astype(np.int32)
    obs_pt = np.repeat(np.arange(M), n)
    X = np.stack([first * 0.2 + rng.uniform(-2, 2, M), rng.uniform(-2, 2, M), rng.uniform(6, 14, M)], 1)
    ph = np.einsum('mck,mk->mc', projs[obs_view][:, :, 0:3], X[obs_pt]) + projs[obs_view][:, :, 3]
    uv = ph[:, 0:2] / ph[:, 2:3] + rng.normal(0, 0.5, (rowptr[-1], 2))
    return projs, rowptr, obs_view, uv, X + rng.normal(0, 0.05, X.shape)


def pnp_problem(P: int, rng):
    """P views with 80..120 observations each.  Returns (rowptr (P+1,) i32, xy (m,2) normalised observations,
    X (m,3) world points, rots (P,3,3) and locs (P,3): the perturbed starting poses)."""
    from scipy.spatial.transform import Rotation
    n = rng.integers(80, 121, P)
    rowptr = np.zeros(P + 1, dtype=np.int32)
    rowptr[1:] = np.cumsum(n)
    m = int(rowptr[-1])
    R = Rotation.from_rotvec(rng.normal(0, 0.2, (P, 3))).as_matrix()
    C = rng.normal(0, 1, (P, 3))
    pc = np.stack([rng.uniform(-3, 3, m), rng.uniform(-2, 2, m), rng.uniform(5, 15, m)], 1)
    view = np.repeat(np.arange(P), n)
    X = np.einsum('mij,mj->mi', R[view], pc) + C[view]
    xy = pc[:, 0:2] / pc[:, 2:3] + rng.normal(0, 1e-3, (m, 2))
    rots = np.einsum('pij,pjk->pik', Rotation.from_rotvec(rng.normal(0, 0.01, (P, 3))).as_matrix(), R)
    locs = C + rng.normal(0, 0.05, (P, 3))
    return rowptr, xy, X, rots, locs
