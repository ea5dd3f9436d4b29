#!/usr/bin/env python3
"""A small tour of the library: every solver mode and entry point once, on small inputs (the command to put under
compute-sanitizer or a debugger; compute-sanitizer is closed on this pool of GPUs).
    python profiles/tour.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sfm_b200                      # noqa: E402
from sfm_b200 import _capi, synthetic  # noqa: E402
from sfm_b200.engine import BAProblem  # noqa: E402

eng = sfm_b200.Engine(device=0)
prob, _ = synthetic.make_shape('small')
print('small', eng.ba_solve_host(prob, lam=5.0, iterations=2).final_cost)
prob, _ = synthetic.make_shape('small')
print('huber', eng.ba_solve_host(prob, lam=1e-3, iterations=3, mode=_capi.MODE_ADAPTIVE_LM, dvdc_sign_quirk=False,
                                 huber_delta=0.002).final_cost)
prob, _ = synthetic.make_shape('tiny')
print('dense', eng.ba_solve_host(prob, lam=5.0, iterations=2, solver=_capi.SOLVER_DENSE_CHOLESKY).final_cost)
# long tracks (multi-chunk groups) and unobserved points
p0, _ = synthetic.make_problem(200, 12, 12 * 200, seed=3)
pts = np.concatenate([[[300.0, 0, 10.0]], p0.pts, [[301.0, 0, 10.0]]])
track = np.concatenate([[0], np.diff(p0.rowptr), [0]])
rowptr = np.zeros(pts.shape[0] + 1, dtype=np.int32)
rowptr[1:] = np.cumsum(track)
obs_pt = np.repeat(np.arange(pts.shape[0]), track).astype(np.int32)
p = BAProblem(cams=p0.cams, pts=pts, obs_uv=p0.obs_uv, obs_cam=p0.obs_cam, obs_pt=obs_pt, rowptr=rowptr)
print('long tracks', eng.ba_solve_host(p, lam=5.0, iterations=2).final_cost)
# refiners and RANSAC
rng = np.random.default_rng(0)
n, H = 500, 64
k = np.array([[800.0, 0, 320], [0, 800, 240], [0, 0, 1]])
X = np.vstack([rng.uniform(-2, 2, (2, n)), rng.uniform(4, 9, (1, n)), np.ones((1, n))])
pp = k @ np.hstack((np.eye(3), np.zeros((3, 1)))) @ X
key = pp / pp[2]
samples = np.array([rng.choice(n, 6, replace=False) for _ in range(H)], dtype=np.int32)
print('pnp ransac', eng.pnp_ransac(key.T, X.T, k, samples, 1.0)['count'])
pairs = rng.uniform(-1, 1, (n, 4))
samples8 = np.array([rng.choice(n, 8, replace=False) for _ in range(H)], dtype=np.int32)
print('fund ransac', eng.fund_ransac(pairs, samples8, 1e-2)['count'])
eng.close()
