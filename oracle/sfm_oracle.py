"""CPU oracle for the bundle-adjustment / nonlinear-refinement hot path.

TEST INFRASTRUCTURE -- NOT PRODUCT CODE.  Only `tests/`, `__graft_entry__.smoke()` and
`bench.py`'s `cpu_baseline` / `--impl reference` legs may import this module, and there only as
the checker or as the timed CPU baseline.  The product path (`sfm_b200`) never imports it and
fails loudly when its CUDA library is missing.

This is a vectorised NumPy (+ scipy.sparse for the large shapes) *restatement* of the reference's
algorithm -- written from the maths, not copied -- with every function citing the reference
file:line it follows.  Parity is PINNED: `tests/test_oracle_golden.py` checks it against fixtures
produced by running the unmodified reference (`tests/golden/make_golden.py`): the four upenn BA
calls, the pipeline's triangulation / PnP calls, the OpenCV PnP known answer, the inline
triangulation known answer and small random BA problems pushed through the literal code.

Conventions: world->camera is p = R^T (X - C); quaternion order (w,x,y,z);
pose vector [Cx,Cy,Cz,qw,qx,qy,qz]; observations are point-major / view-minor
(ba_processor.py:304-306).

Three reference behaviours are reproduced on purpose (flags default to the reference's):
  * dv/dC sign quirk            campose_processor.py:802-804
  * PnP row-packing quirk       campose_processor.py:404-405
  * is_visible index-0 quirk    key_tracker.py:198-204
"""
from __future__ import annotations

import math

import numpy as np

__all__ = [
    'rotation_to_quaternion', 'quaternion_to_rotation', 'verify_rotation_mat',
    'is_visible_literal', 'build_visibility', 'normalise_observations',
    'quaternion_jacobian', 'residual_jacobian', 'ba_cost', 'ba_solve', 'ba_normal_equations',
    'nonlinear_triangulate', 'triangulation_jacobian', 'nonlinear_pnp', 'pose_jacobian',
    'huber_rho', 'huber_weight', 'pnp_dlt_six', 'pnp_ransac', 'fund_eight', 'fund_ransac', 'fund_normalize', 'fund_denormalize',
]


# ----------------------------------------------------------------------------------------------
# a20 / a6: quaternion <-> rotation                                           utils.py:28-105
# ----------------------------------------------------------------------------------------------
def verify_rotation_mat(rot: np.ndarray) -> bool:
    """utils.py:101-105 -- one-sided 1e-8 tolerance on det and on inv(R) - R^T."""
    rot = np.asarray(rot)
    if rot.shape != (3, 3):
        return False
    if np.linalg.det(rot) - 1 >= 1e-8:
        return False
    if np.any((np.linalg.inv(rot) - rot.T) > 1e-8):
        return False
    return True


def rotation_to_quaternion(rot: np.ndarray, check: bool = True) -> np.ndarray:
    """utils.py:28-60.  qw > 0 branch only; raises if R is invalid or |qw| < 1e-6.
    Accepts (3,3) or (V,3,3); returns (4,) or (V,4) in (w,x,y,z) order."""
    rot = np.asarray(rot, dtype=np.float64)
    single = rot.ndim == 2
    r = rot.reshape(-1, 3, 3)
    if check:
        for m in r:
            if not verify_rotation_mat(m):
                raise ValueError('rotation_to_quaternion : Invalid input rotation matrix \n {}'.format(m))
    qw = np.sqrt(1 + r[:, 0, 0] + r[:, 1, 1] + r[:, 2, 2]) / 2.0
    if np.any(np.abs(qw) < 1e-6):
        raise ValueError('rotation_to_quaternion : Invalid output qw')
    qx = (r[:, 2, 1] - r[:, 1, 2]) / (4 * qw)
    qy = (r[:, 0, 2] - r[:, 2, 0]) / (4 * qw)
    qz = (r[:, 1, 0] - r[:, 0, 1]) / (4 * qw)
    q = np.stack([qw, qx, qy, qz], axis=1)
    return q[0] if single else q


def quaternion_to_rotation(q: np.ndarray, check: bool = False) -> np.ndarray:
    """utils.py:64-97.  (4,) -> (3,3) or (V,4) -> (V,3,3).  `check` reproduces the validity raise."""
    q = np.asarray(q, dtype=np.float64)
    single = q.ndim == 1
    q = q.reshape(-1, 4)
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
    if check:
        for m in r:
            if not verify_rotation_mat(m):
                raise ValueError('quaternion_to_rotation : Invalid output rotation matrix \n {}'.format(m))
    return r[0] if single else r


# ----------------------------------------------------------------------------------------------
# a3 / a4: visibility index                                    key_tracker.py:29-59, 198-204
# ----------------------------------------------------------------------------------------------
def is_visible_literal(diag_row: np.ndarray, tri_pt_idx: int) -> int:
    """key_tracker.py:198-204 word for word in behaviour: np.any() is applied to the *indices*
    returned by np.where, so a point whose only matching key index is 0 is reported invisible."""
    key_idx = np.where(diag_row == tri_pt_idx)
    if np.any(key_idx):
        return int(key_idx[0][0])
    return -1


def build_visibility(diag_rows, num_pts: int):
    """Observation list in the loop order of ba_processor.py:304-310 (point-major, view-minor).

    diag_rows[v] is `track_list[v].table[v, :]` (int, -1 = unused).  Returns int64 arrays
    (obs_pt, obs_cam, obs_key).  Vectorised; `tests` check it against `is_visible_literal`."""
    pts, cams, keys = [], [], []
    for v, row in enumerate(diag_rows):
        row = np.asarray(row)
        idx = np.nonzero((row >= 0) & (row < num_pts))[0]
        if idx.size == 0:
            continue
        vals = row[idx]
        order = np.argsort(vals, kind='stable')           # ties keep ascending key index
        sv, si = vals[order], idx[order]
        first = np.ones(sv.size, dtype=bool)
        first[1:] = sv[1:] != sv[:-1]
        starts = np.nonzero(first)[0]
        counts = np.diff(np.append(starts, sv.size))
        t = sv[starts]
        k = si[starts]                                      # smallest matching key index
        keep = (k > 0) | (counts > 1)                       # np.any(indices) quirk
        pts.append(t[keep])
        cams.append(np.full(int(keep.sum()), v, dtype=np.int64))
        keys.append(k[keep])
    if not pts:
        z = np.zeros(0, dtype=np.int64)
        return z, z.copy(), z.copy()
    pts = np.concatenate(pts).astype(np.int64)
    cams = np.concatenate(cams)
    keys = np.concatenate(keys).astype(np.int64)
    order = np.lexsort((cams, pts))
    return pts[order], cams[order], keys[order]


def normalise_observations(obs_px: np.ndarray, obs_cam: np.ndarray, k_mats: np.ndarray) -> np.ndarray:
    """a5, ba_processor.py:339-342: b = (K^-1 [u,v,1]^T) / z, first two rows.  obs_px (m,2) are the
    float32 cv2 key-point coordinates promoted to float64; k_mats (V,3,3)."""
    kinv = np.linalg.inv(np.asarray(k_mats, dtype=np.float64))
    h = np.concatenate([np.asarray(obs_px, dtype=np.float64), np.ones((obs_px.shape[0], 1))], axis=1)
    c = np.einsum('mab,mb->ma', kinv[obs_cam], h)
    return c[:, 0:2] / c[:, 2:3]


# ----------------------------------------------------------------------------------------------
# a8-a12: per-observation residual and Jacobians
# ----------------------------------------------------------------------------------------------
def quaternion_jacobian(q: np.ndarray) -> np.ndarray:
    """a9, campose_processor.py:636-702: d vec_row(R) / d q, (V,9,4), row 3k+c <-> R[k][c]."""
    q = np.asarray(q, dtype=np.float64).reshape(-1, 4)
    w, x, y, z = (2 * q[:, i] for i in range(4))
    o = np.zeros_like(w)
    rows = [
        [o, o, -2 * y, -2 * z],
        [-z, y, x, -w],
        [y, z, w, x],
        [z, y, x, w],
        [o, -2 * x, o, -2 * z],
        [-x, -w, z, y],
        [-y, z, -w, x],
        [x, w, z, y],
        [o, -2 * x, -2 * y, o],
    ]
    return np.stack([np.stack(r, axis=1) for r in rows], axis=1)


def residual_jacobian(cams, pts, obs_cam, obs_pt, obs_b, dvdc_quirk: bool = True, want_jac: bool = True):
    """a7 + a8-a12 for every observation.

    cams (V,7) = [C, q]; pts (N,3); obs_b (m,2) normalised observations.
    Follows ba_processor.py:317-349: R = q->R, P = [R^T | R^T(-C)], f = (P X~)/z, r = b - f;
    J_p = campose_processor.construct_jacobian_matrix (462-482, 636-808; q re-derived from R at
    464), J_x = triangulation_processor.construct_jacobian_matrix (237-271) with the K-free P.
    Returns r (m,2), Jp (m,2,7), Jx (m,2,3)."""
    cams = np.asarray(cams, dtype=np.float64)
    pts = np.asarray(pts, dtype=np.float64)
    rot = quaternion_to_rotation(cams[:, 3:7])                       # (V,3,3)
    tvec = np.einsum('vkc,vk->vc', rot, -cams[:, 0:3])               # R^T (-C)
    R = rot[obs_cam]
    X = pts[obs_pt]
    p = np.einsum('mkc,mk->mc', R, X) + tvec[obs_cam]                # proj @ [X;1]
    pu, pv, pw = p[:, 0], p[:, 1], p[:, 2]
    f = p[:, 0:2] / p[:, 2:3]
    r = np.asarray(obs_b, dtype=np.float64) - f
    if not want_jac:
        return r, None, None
    m = p.shape[0]
    pw2 = np.square(pw)
    # a12: (pw * P[0,:3] - pu * P[2,:3]) / pw^2 with P[c,k] = R[k][c]
    jx = np.empty((m, 2, 3))
    jx[:, 0, :] = (pw[:, None] * R[:, :, 0] - pu[:, None] * R[:, :, 2]) / pw2[:, None]
    jx[:, 1, :] = (pw[:, None] * R[:, :, 1] - pv[:, None] * R[:, :, 2]) / pw2[:, None]
    # a11: centre block, v-row sign quirk at campose_processor.py:802-804
    jpc = np.empty((m, 2, 3))
    jpc[:, 0, :] = (pw[:, None] * -R[:, :, 0] - pu[:, None] * -R[:, :, 2]) / pw2[:, None]
    if dvdc_quirk:
        jpc[:, 1, :] = (pw[:, None] * -R[:, :, 1] - pv[:, None] * R[:, :, 2]) / pw2[:, None]
    else:
        jpc[:, 1, :] = (pw[:, None] * -R[:, :, 1] - pv[:, None] * -R[:, :, 2]) / pw2[:, None]
    # a10: d(u,v)/d vec_row(R), d = X - C
    d = X - cams[obs_cam, 0:3]
    jrr = np.zeros((m, 2, 9))
    for k in range(3):
        jrr[:, 0, 3 * k + 0] = pw * d[:, k]
        jrr[:, 0, 3 * k + 2] = -pu * d[:, k]
        jrr[:, 1, 3 * k + 1] = pw * d[:, k]
        jrr[:, 1, 3 * k + 2] = -pv * d[:, k]
    jrr /= pw2[:, None, None]
    # a9 with q re-derived from R (qw > 0 representative), campose_processor.py:464
    qj = rotation_to_quaternion(rot, check=False)
    jrq = quaternion_jacobian(qj)
    jpq = np.einsum('mab,mbc->mac', jrr, jrq[obs_cam])
    jp = np.concatenate([jpc, jpq], axis=2)
    return r, jp, jx


def pose_jacobian(rot, loc, pt_3d, dvdc_quirk: bool = True) -> np.ndarray:
    """CamposeProcessor.construct_jacobian_matrix(rot, loc, pt_3d) (campose_processor.py:462-482)
    for a single observation; rot (3,3) is used as given, q is re-derived from it."""
    rot = np.asarray(rot, dtype=np.float64)
    loc = np.asarray(loc, dtype=np.float64).reshape(3)
    X = np.asarray(pt_3d, dtype=np.float64).reshape(-1)[0:3]
    q = rotation_to_quaternion(rot, check=True)
    p = rot.T @ X + rot.T @ -loc
    pu, pv, pw = p
    d = X - loc
    jrr = np.zeros((2, 9))
    for k in range(3):
        jrr[0, 3 * k + 0] = pw * d[k]
        jrr[0, 3 * k + 2] = -pu * d[k]
        jrr[1, 3 * k + 1] = pw * d[k]
        jrr[1, 3 * k + 2] = -pv * d[k]
    jrr /= pw * pw
    jpq = jrr @ quaternion_jacobian(q)[0]
    jpc = np.empty((2, 3))
    jpc[0] = (pw * -rot[:, 0] - pu * -rot[:, 2]) / (pw * pw)
    sgn = 1.0 if dvdc_quirk else -1.0
    jpc[1] = (pw * -rot[:, 1] - pv * sgn * rot[:, 2]) / (pw * pw)
    return np.hstack((jpc, jpq))


def ba_cost(cams, pts, obs_cam, obs_pt, obs_b) -> float:
    """Sum ||b - f||^2 in normalised camera coordinates (the known-answer quantity of SURVEY 8(c))."""
    r, _, _ = residual_jacobian(cams, pts, obs_cam, obs_pt, obs_b, want_jac=False)
    return float(np.sum(np.square(r)))


# ----------------------------------------------------------------------------------------------
# a13-a16: normal equations, Schur complement, update            ba_processor.py:355-406
# ----------------------------------------------------------------------------------------------
def _segment_sum(values: np.ndarray, seg: np.ndarray, nseg: int) -> np.ndarray:
    """Sum `values` (m, ...) into `nseg` rows by integer key `seg` (any order)."""
    flat = values.reshape(values.shape[0], -1)
    out = np.zeros((nseg, flat.shape[1]))
    if flat.shape[0]:
        order = np.argsort(seg, kind='stable')
        s = seg[order]
        starts = np.nonzero(np.r_[True, s[1:] != s[:-1]])[0]
        out[s[starts]] = np.add.reduceat(flat[order], starts, axis=0)
    return out.reshape((nseg,) + values.shape[1:])


def huber_rho(s: np.ndarray, delta: float) -> np.ndarray:
    """Huber loss of the squared residual norm s = |r|^2: s inside delta, 2 delta sqrt(s) - delta^2 outside
    (not in the reference: SURVEY.md section 8(f) row 4)."""
    s = np.asarray(s, dtype=np.float64)
    return np.where(s <= delta * delta, s, 2.0 * delta * np.sqrt(np.maximum(s, 1e-300)) - delta * delta)


def huber_weight(s: np.ndarray, delta: float) -> np.ndarray:
    """rho'(s): 1 inside delta, delta / sqrt(s) outside -- the IRLS weight of one observation."""
    s = np.asarray(s, dtype=np.float64)
    return np.where(s <= delta * delta, 1.0, delta / np.sqrt(np.maximum(s, 1e-300)))


def ba_normal_equations(cams, pts, obs_cam, obs_pt, obs_b, lam: float, dvdc_quirk: bool = True,
                        huber_delta: float = 0.0):
    """One linearisation: returns dict with r, Jp, Jx, A (V,7,7), e_p (V,7), D (N,3,3), Dinv,
    e_x (N,3), W (m,7,3) following ba_processor.py:355-379 block by block.  huber_delta > 0: every
    observation's residual and Jacobian rows are scaled by sqrt(rho'(|r|^2)) first (iteratively
    re-weighted least squares; r in the result stays unweighted, `cost` is sum rho)."""
    V, N = cams.shape[0], pts.shape[0]
    r, jp, jx = residual_jacobian(cams, pts, obs_cam, obs_pt, obs_b, dvdc_quirk)
    r_raw = r
    if huber_delta > 0:
        sw = np.sqrt(huber_weight(np.sum(r * r, axis=1), huber_delta))
        r = r * sw[:, None]
        jp = jp * sw[:, None, None]
        jx = jx * sw[:, None, None]
    D = _segment_sum(np.einsum('mai,maj->mij', jx, jx), obs_pt, N) + lam * np.eye(3)      # a13
    e_x = _segment_sum(np.einsum('mai,ma->mi', jx, r), obs_pt, N)
    A = _segment_sum(np.einsum('mai,maj->mij', jp, jp), obs_cam, V) + lam * np.eye(7)     # a14
    e_p = _segment_sum(np.einsum('mai,ma->mi', jp, r), obs_cam, V)
    W = np.einsum('mai,maj->mij', jp, jx)                                                  # B blocks
    Dinv = np.linalg.inv(D)
    s_raw = np.sum(r_raw * r_raw, axis=1)
    cost = float(np.sum(huber_rho(s_raw, huber_delta))) if huber_delta > 0 else float(np.sum(s_raw))
    return dict(r=r_raw, jp=jp, jx=jx, A=A, e_p=e_p, D=D, Dinv=Dinv, e_x=e_x, W=W, cost=cost)


def _reduced_system_dense(ne, obs_cam, obs_pt, V):
    """S = A - B D^-1 B^T and g = e_p - B D^-1 e_x assembled densely (a15, ba_processor.py:382)."""
    W, Dinv = ne['W'], ne['Dinv']
    m = W.shape[0]
    S = np.zeros((7 * V, 7 * V))
    for i in range(V):
        S[7 * i:7 * i + 7, 7 * i:7 * i + 7] = ne['A'][i]
    E = np.einsum('mij,mjk->mik', W, Dinv[obs_pt])                    # W_o D_j^-1, (m,7,3)
    g = ne['e_p'] - _segment_sum(np.einsum('mij,mj->mi', E, ne['e_x'][obs_pt]), obs_cam, V)
    # pairs of observations of the same point
    order = np.argsort(obs_pt, kind='stable')
    sp = obs_pt[order]
    starts = np.nonzero(np.r_[True, sp[1:] != sp[:-1]])[0] if m else np.zeros(0, dtype=int)
    ends = np.r_[starts[1:], m] if m else starts
    lens = ends - starts
    for L in np.unique(lens):
        sel = starts[lens == L]
        idx = order[sel[:, None] + np.arange(L)[None, :]]             # (P, L) observation ids
        for a in range(L):
            for b in range(L):
                blocks = np.einsum('pij,pkj->pik', E[idx[:, a]], W[idx[:, b]])
                ca, cb = obs_cam[idx[:, a]], obs_cam[idx[:, b]]
                rows = (7 * ca[:, None, None] + np.arange(7)[None, :, None])
                cols = (7 * cb[:, None, None] + np.arange(7)[None, None, :])
                np.subtract.at(S, (np.broadcast_to(rows, blocks.shape), np.broadcast_to(cols, blocks.shape)), blocks)
    return S, g.reshape(-1)


def _reduced_system_sparse(ne, obs_cam, obs_pt, V, N):
    """Same reduced system through scipy.sparse BSR products, for BAL-sized shapes where the
    pair loop above and the reference's dense (7V x 3N) B cannot be formed."""
    import scipy.sparse as sp
    W, Dinv = ne['W'], ne['Dinv']
    m = W.shape[0]
    order = np.lexsort((obs_pt, obs_cam))
    indptr = np.searchsorted(obs_cam[order], np.arange(V + 1))
    B = sp.bsr_matrix((W[order], obs_pt[order], indptr), shape=(7 * V, 3 * N), blocksize=(7, 3))
    Di = sp.bsr_matrix((Dinv, np.arange(N), np.arange(N + 1)), shape=(3 * N, 3 * N), blocksize=(3, 3))
    BD = (B @ Di).tocsr()
    S = sp.bsr_matrix((ne['A'], np.arange(V), np.arange(V + 1)), shape=(7 * V, 7 * V)).tocsr() - BD @ B.T.tocsr()
    g = ne['e_p'].reshape(-1) - BD @ ne['e_x'].reshape(-1)
    return S, g, B


def ba_solve(cams, pts, obs_cam, obs_pt, obs_b, lam: float = 5.0, iterations: int = 3,
             dvdc_quirk: bool = True, sparse: bool | None = None, return_costs: bool = True,
             huber_delta: float = 0.0):
    """BaProcessor.__execute_bundle_adjustment in reference mode (ba_processor.py:297-406):
    fixed damping `lam`, `iterations` always-accepted steps, reduced system solved exactly,
    additive quaternion update followed by normalisation (383-392), dx back-substitution (405-406).

    Returns (cams, pts, costs) with costs[k] = cost at the start of iteration k and costs[-1] the
    final cost."""
    cams = np.array(cams, dtype=np.float64, copy=True)
    pts = np.array(pts, dtype=np.float64, copy=True)
    obs_cam = np.asarray(obs_cam, dtype=np.int64)
    obs_pt = np.asarray(obs_pt, dtype=np.int64)
    V, N = cams.shape[0], pts.shape[0]
    if sparse is None:
        sparse = obs_cam.shape[0] > 20000
    costs = []
    for _ in range(iterations):
        ne = ba_normal_equations(cams, pts, obs_cam, obs_pt, obs_b, lam, dvdc_quirk, huber_delta)
        costs.append(ne['cost'])
        if sparse:
            S, g, _ = _reduced_system_sparse(ne, obs_cam, obs_pt, V, N)
            dp = np.linalg.solve(S.toarray(), g)
        else:
            S, g = _reduced_system_dense(ne, obs_cam, obs_pt, V)
            dp = np.linalg.solve(S, g)
        dp = dp.reshape(V, 7)
        # dx = D^-1 (e_x - B^T dp)
        bt_dp = _segment_sum(np.einsum('mij,mi->mj', ne['W'], dp[obs_cam]), obs_pt, N)
        dx = np.einsum('nij,nj->ni', ne['Dinv'], ne['e_x'] - bt_dp)
        cams += dp
        cams[:, 3:7] /= np.sqrt(np.sum(np.square(cams[:, 3:7]), axis=1, keepdims=True))
        pts += dx
    if return_costs:
        if huber_delta > 0:
            rr, _, _ = residual_jacobian(cams, pts, obs_cam, obs_pt, obs_b, want_jac=False)
            costs.append(float(np.sum(huber_rho(np.sum(rr * rr, axis=1), huber_delta))))
        else:
            costs.append(ba_cost(cams, pts, obs_cam, obs_pt, obs_b))
    return cams, pts, costs


# ----------------------------------------------------------------------------------------------
# a18: nonlinear triangulation                        triangulation_processor.py:160-309
# ----------------------------------------------------------------------------------------------
def triangulation_jacobian(X, projs):
    """triangulation_processor.py:237-271 for a batch: X (M,3), projs (n,3,4) -> (M,2n,3) and the
    homogeneous projections (M,n,3)."""
    X = np.asarray(X, dtype=np.float64)
    projs = np.asarray(projs, dtype=np.float64)
    p = np.einsum('nck,mk->mnc', projs[:, :, 0:3], X) + projs[None, :, :, 3]
    pw2 = np.square(p[:, :, 2])
    ju = (p[:, :, 2, None] * projs[None, :, 0, 0:3] - p[:, :, 0, None] * projs[None, :, 2, 0:3]) / pw2[:, :, None]
    jv = (p[:, :, 2, None] * projs[None, :, 1, 0:3] - p[:, :, 1, None] * projs[None, :, 2, 0:3]) / pw2[:, :, None]
    jac = np.stack([ju, jv], axis=2).reshape(X.shape[0], -1, 3)
    return jac, p


def nonlinear_triangulate(init_3d_pts, projs, matched_pairs, damping_factor: float = 0.5, iteration: int = 100):
    """TriangulationProcessor.nonlinear_triangulate (triangulation_processor.py:160-234):
    X -= inv(J^T J + lam I3) J^T (proj - obs), `iteration` times, every point independent.
    init_3d_pts (4,M); projs list of n (3,4) (K included by the caller); matched_pairs list of n
    (3,M).  Returns a new (4,M) array; the input is not mutated (np.copy at :208)."""
    out = np.array(init_3d_pts, dtype=np.float64, copy=True)
    projs = np.stack([np.asarray(p, dtype=np.float64) for p in projs])
    obs = np.stack([np.asarray(mp, dtype=np.float64)[0:2, :].T for mp in matched_pairs], axis=1)  # (M,n,2)
    X = out[0:3, :].T.copy()
    eye = damping_factor * np.eye(3)
    for _ in range(iteration):
        jac, p = triangulation_jacobian(X, projs)
        err = (p[:, :, 0:2] / p[:, :, 2:3] - obs).reshape(X.shape[0], -1)          # (M,2n)
        H = np.einsum('mai,maj->mij', jac, jac) + eye
        g = np.einsum('mai,ma->mi', jac, err)
        X = X - np.einsum('mij,mj->mi', np.linalg.inv(H), g)
    out[0:3, :] = X.T
    return out


# ----------------------------------------------------------------------------------------------
# f2: linear triangulation + cheirality, the step before a18
#                                triangulation_processor.py:91-157, campose_processor.py:139-189
# ----------------------------------------------------------------------------------------------
def linear_triangulate(projs, matched_pairs):
    """TriangulationProcessor.linear_triangulate (triangulation_processor.py:136-155) for a batch:
    rows u P[2] - P[0] and v P[2] - P[1] per view, the right singular vector of the smallest singular
    value, divided by its last entry.  projs list of n (3,4); matched_pairs list of n (3,M).
    Returns (4,M).  (The reference accepts exactly two views, :122-134; n views follow the same rows.)"""
    projs = np.stack([np.asarray(p, dtype=np.float64) for p in projs])                       # (n,3,4)
    uv = np.stack([np.asarray(mp, dtype=np.float64)[0:2, :].T for mp in matched_pairs], axis=1)  # (M,n,2)
    M, n = uv.shape[0], uv.shape[1]
    a = np.empty((M, 2 * n, 4))
    a[:, 0::2, :] = uv[:, :, 0, None] * projs[None, :, 2, :] - projs[None, :, 0, :]
    a[:, 1::2, :] = uv[:, :, 1, None] * projs[None, :, 2, :] - projs[None, :, 1, :]
    _, _, vh = np.linalg.svd(a)
    x = vh[:, -1, :]
    return (x / x[:, 3:4]).T


def cheirality_indices(p1, p2, tri_3d_pts):
    """CamposeProcessor.evalulate_cam_pose_cheirality (campose_processor.py:181-189): indices of the
    points whose third projected coordinate is positive in both views."""
    X = np.asarray(tri_3d_pts, dtype=np.float64)
    d1 = (np.asarray(p1, dtype=np.float64) @ X)[2]
    d2 = (np.asarray(p2, dtype=np.float64) @ X)[2]
    return [int(i) for i in np.flatnonzero((d1 > 0) & (d2 > 0))]


# ----------------------------------------------------------------------------------------------
# a19: nonlinear PnP                                           campose_processor.py:308-459
# ----------------------------------------------------------------------------------------------
def nonlinear_pnp(key_2d_pts, tri_3d_pts, intrinsic_mat, init_rot, init_loc, damping_factor: float,
                  iteration: int, row_packing_quirk: bool = True, dvdc_quirk: bool = True):
    """CamposeProcessor.nonlinear_estimate_cam_pose_pnp.  key_2d_pts (3,n) pixels, tri_3d_pts (4,n).

    Row-packing quirk (campose_processor.py:404-405): `jac_rp_all[pt_idx:pt_idx+2]` makes rows
    0..n-1 the u-rows of points 0..n-1, row n the v-row of the last point, the rest zero.
    Iteration 0 uses init_rot as given (367, 388, 397); later ones use q->R of the normalised q."""
    key = np.asarray(key_2d_pts, dtype=np.float64)
    tri = np.asarray(tri_3d_pts, dtype=np.float64)
    n = key.shape[1]
    if n != tri.shape[1]:
        raise ValueError('key pts num - triangulated pts num : {} - {}'.format(n, tri.shape[1]))
    kinv = np.linalg.inv(np.asarray(intrinsic_mat, dtype=np.float64))
    cam2d = kinv @ key
    b = (cam2d[0:2] / cam2d[2:3]).T                                            # (n,2)
    X = tri[0:3].T
    w4 = tri[3]
    q = rotation_to_quaternion(np.asarray(init_rot, dtype=np.float64))
    q = q / math.sqrt(np.sum(np.square(q)))
    loc = np.array(init_loc, dtype=np.float64).reshape(3).copy()
    rot = np.array(init_rot, dtype=np.float64, copy=True)
    for _ in range(iteration):
        qj = rotation_to_quaternion(rot, check=False)
        jrq = quaternion_jacobian(qj)[0]
        p = X @ rot + w4[:, None] * (rot.T @ -loc)[None, :]
        pu, pv, pw = p[:, 0], p[:, 1], p[:, 2]
        err = b - p[:, 0:2] / p[:, 2:3]
        pw2 = pw * pw
        d = X - loc[None, :]
        jrr = np.zeros((n, 2, 9))
        for k in range(3):
            jrr[:, 0, 3 * k + 0] = pw * d[:, k]
            jrr[:, 0, 3 * k + 2] = -pu * d[:, k]
            jrr[:, 1, 3 * k + 1] = pw * d[:, k]
            jrr[:, 1, 3 * k + 2] = -pv * d[:, k]
        jrr /= pw2[:, None, None]
        jpq = jrr @ jrq
        jpc = np.empty((n, 2, 3))
        jpc[:, 0, :] = (pw[:, None] * -rot[None, :, 0] - pu[:, None] * -rot[None, :, 2]) / pw2[:, None]
        sgn = 1.0 if dvdc_quirk else -1.0
        jpc[:, 1, :] = (pw[:, None] * -rot[None, :, 1] - pv[:, None] * sgn * rot[None, :, 2]) / pw2[:, None]
        jp = np.concatenate([jpc, jpq], axis=2)                                # (n,2,7)
        if row_packing_quirk:
            J = np.concatenate([jp[:, 0, :], jp[n - 1:n, 1, :]], axis=0)
            e = np.concatenate([err[:, 0], err[n - 1:n, 1]])
        else:
            J = jp.reshape(2 * n, 7)
            e = err.reshape(2 * n)
        delta = np.linalg.inv(J.T @ J + damping_factor * np.identity(7)) @ (J.T @ e)
        loc = loc + delta[0:3]
        q = q + delta[3:7]
        q = q / math.sqrt(np.sum(np.square(q)))
        rot = quaternion_to_rotation(q, check=True)
    rot = quaternion_to_rotation(q, check=True)
    return rot, loc.reshape(3, 1)


# ----------------------------------------------------------------------------------------------
# f3: RANSAC initialisers        campose_processor.py:485-633, epipolar_processor.py:96-265
# The sample index table is an input (the reference draws it with random.sample, one draw of 6 / 8
# indices per hypothesis, campose_processor.py:531 / epipolar_processor.py:222), so the oracle and
# the device code see the reference's draws.
# ----------------------------------------------------------------------------------------------
def pnp_dlt_six(p2d: np.ndarray, p3d: np.ndarray):
    """Pose from six correspondences (campose_processor.py:565-633).  p2d (3,6) camera coordinates,
    p3d (4,6).  Returns (rot (3,3), loc (3,1), lapack_sign) where lapack_sign = +1 when numpy's SVD
    returned the last right singular vector with det(rot_t_noise) > 0, else -1: in the second case
    lines 628-631 flip rot AND loc, which leaves (R, -C) instead of (R, C)."""
    w = np.zeros((12, 12))
    for i in range(6):
        x, y, z = p2d[0, i], p2d[1, i], p2d[2, i]
        X = p3d[0:3, i]
        w[2 * i, 0:3] = z * X
        w[2 * i, 3] = z                                    # :592 (4th column is p2d_z, not p2d_z * W)
        w[2 * i, 8:11] = -x * X
        w[2 * i, 11] = -x
        w[2 * i + 1, 4:7] = z * X
        w[2 * i + 1, 7] = z
        w[2 * i + 1, 8:11] = -y * X
        w[2 * i + 1, 11] = -y
    _, _, vh = np.linalg.svd(w)                            # :616
    cam_mat = vh[-1, :].reshape(3, 4)
    uu, ss, vvh = np.linalg.svd(cam_mat[:, 0:3])           # :621
    rot = (uu @ vvh).T
    loc = (rot @ -cam_mat[:, 3:4]) / ss[0]                 # :625
    sign = 1
    if np.linalg.det(rot) < 0:                             # :628-631
        rot = -rot
        loc = -loc
        sign = -1
    return rot, loc, sign


def pnp_ransac(key_2d_pts, tri_3d_pts, intrinsic_mat, samples, inlier_threshold: float):
    """campose_processor.py:485-563 for a given sample table (H,6).  Returns a dict with the per-hypothesis
    models (H,12: rot row-major | loc), LAPACK signs, inlier counts, the winner, its inlier list and pose."""
    key = np.asarray(key_2d_pts, dtype=float)
    tri = np.asarray(tri_3d_pts, dtype=float)
    k = np.asarray(intrinsic_mat, dtype=float)
    kinv = np.linalg.inv(k)
    H = len(samples)
    models = np.zeros((H, 12))
    signs = np.zeros(H, dtype=np.int32)
    counts = np.zeros(H, dtype=np.int32)
    best = dict(hyp=-1, count=0, inliers=[], rot=np.identity(3), loc=np.zeros((3, 1)))
    for h in range(H):
        idx = list(samples[h])
        rot, loc, sg = pnp_dlt_six(kinv @ key[:, idx], tri[:, idx])
        proj = k @ np.hstack((rot.T, rot.T @ -loc))                      # :540
        p = proj @ tri
        with np.errstate(all='ignore'):
            p = p / p[2]
            err = np.sqrt(np.sum((key - p) ** 2, axis=0))                # :551
        inl = np.flatnonzero(err < inlier_threshold)
        models[h, 0:9] = rot.reshape(9)
        models[h, 9:12] = loc.reshape(3)
        signs[h] = sg
        counts[h] = inl.size
        if inl.size > best['count']:                                     # :558
            best = dict(hyp=h, count=int(inl.size), inliers=[int(i) for i in inl], rot=rot, loc=loc)
    best.update(models=models, signs=signs, counts=counts)
    return best


def fund_normalize(left_xy: np.ndarray, right_xy: np.ndarray):
    """epipolar_processor.py:96-137.  left_xy / right_xy (n,2).  Returns (pairs (n,4), T_left, T_right)."""
    n = left_xy.shape[0]
    lm, rm = np.average(left_xy, axis=0), np.average(right_xy, axis=0)
    ls = (2 * n) ** 0.5 / np.sum(np.sum((left_xy - lm) ** 2, axis=1) ** 0.5)
    rs = (2 * n) ** 0.5 / np.sum(np.sum((right_xy - rm) ** 2, axis=1) ** 0.5)
    tl = np.array([[ls, 0, -lm[0] * ls], [0, ls, -lm[1] * ls], [0, 0, 1]])
    tr = np.array([[rs, 0, -rm[0] * rs], [0, rs, -rm[1] * rs], [0, 0, 1]])
    one = np.ones((n, 1))
    ln = (tl @ np.column_stack((left_xy, one)).T).T
    rn = (tr @ np.column_stack((right_xy, one)).T).T
    return np.column_stack((ln[:, 0:2], rn[:, 0:2])), tl, tr


def fund_denormalize(fund_mat, t_left, t_right):
    """epipolar_processor.py:250-265."""
    f = t_right.T @ fund_mat @ t_left
    return f / f[2][2]


def fund_eight(pairs8: np.ndarray) -> np.ndarray:
    """Eight-point fundamental matrix with the rank-2 projection (epipolar_processor.py:140-191)."""
    x1, y1, x1_, y1_ = pairs8[:, 0], pairs8[:, 1], pairs8[:, 2], pairs8[:, 3]
    w = np.column_stack((x1 * x1_, y1 * x1_, x1_, x1 * y1_, y1 * y1_, y1_, x1, y1, np.ones(8)))
    _, _, vh = np.linalg.svd(w)
    f_ = vh[8, :].reshape(3, 3)
    u, s, vh2 = np.linalg.svd(f_)
    f2 = u @ np.diag([s[0], s[1], 0.0]) @ vh2
    if np.linalg.matrix_rank(f2) != 2:                                    # :179-183
        raise ValueError('f__ rank is not equal to 2')
    return f2 / f2[2][2]


def fund_ransac(pairs: np.ndarray, samples, inlier_threshold: float):
    """epipolar_processor.py:194-247 (the rows > 8 branch) for a given sample table (H,8)."""
    pairs = np.asarray(pairs, dtype=float)
    H = len(samples)
    models = np.zeros((H, 9))
    counts = np.zeros(H, dtype=np.int32)
    best = dict(hyp=-1, count=0, inliers=None, F=np.zeros((3, 3)))
    xl = np.column_stack((pairs[:, 0:2], np.ones(pairs.shape[0])))
    xr = np.column_stack((pairs[:, 2:4], np.ones(pairs.shape[0])))
    for h in range(H):
        f = fund_eight(pairs[list(samples[h]), :])
        val = np.abs(np.sum(xr * (f @ xl.T).T, axis=1))                   # :235
        inl = np.flatnonzero(val < inlier_threshold)
        models[h] = f.reshape(9)
        counts[h] = inl.size
        if inl.size > best['count']:                                      # :239
            best = dict(hyp=h, count=int(inl.size), inliers=[int(i) for i in inl], F=f)
    best.update(models=models, counts=counts)
    return best
