#!/usr/bin/env python3
"""Config C3 of SURVEY.md section 8 (batched refiners) and the other BA shapes, one JSON line each.

    python profiles/bench_refiners.py [--points 1000000] [--views 10000]

* a18 batched: M points with 2..6 observations each out of 200 projections, `iterations` damped steps
  (reference defaults: lambda 0.5, 100 iterations, triangulation_processor.py:12)
* a19 batched: P views with ~100 observations each (reference pipeline: lambda 5, 300 iterations)
* f3: both RANSAC initialisers, 100 000 points x 2 000 hypotheses
* BA at the Ladybug and Final shapes (device-resident, reference mode), as sanity lines next to the
  Venice headline of bench.py.
Times are wall-clock around the host-buffer entry points (uploads and downloads included) and, for the
BA shapes, CUDA-event device time as reported by the library.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sfm_b200                      # noqa: E402
from sfm_b200 import synthetic       # noqa: E402


tri_problem, pnp_problem = synthetic.tri_problem, synthetic.pnp_problem


def timed(fn, reps=3):
    fn()
    best = 1e30
    for _ in range(reps):
        t0 = time.perf_counter()
        out = fn()
        best = min(best, time.perf_counter() - t0)
    return best, out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--points', type=int, default=1000000)
    ap.add_argument('--views', type=int, default=10000)
    ap.add_argument('--tri-iterations', type=int, default=100)
    ap.add_argument('--pnp-iterations', type=int, default=300)
    ap.add_argument('--skip-ba', action='store_true')
    args = ap.parse_args()
    eng = sfm_b200.Engine(device=0)
    rng = np.random.default_rng(3)

    projs, rowptr, obs_view, uv, X0 = tri_problem(args.points, rng)
    dt, out = timed(lambda: eng.triangulate_refine(projs, rowptr, obs_view, uv, X0, 0.5, args.tri_iterations))
    print(json.dumps({'workload': 'C3 a18 batched nonlinear triangulation', 'points': args.points,
                      'observations': int(rowptr[-1]), 'iterations': args.tri_iterations, 'seconds': dt,
                      'point_iterations_per_sec': args.points * args.tri_iterations / dt,
                      'mean_abs_move': float(np.mean(np.abs(out - X0)))}))

    dt, (lin, front) = timed(lambda: eng.triangulate_linear(projs, rowptr, obs_view, uv))
    print(json.dumps({'workload': 'f2 batched linear (DLT) triangulation + cheirality', 'points': args.points,
                      'observations': int(rowptr[-1]), 'seconds': dt, 'points_per_sec': args.points / dt,
                      'front_fraction': float(np.mean(front)), 'mean_abs_diff_to_refined': float(np.mean(np.abs(lin - out)))}))

    rowptr, xy, X, rots, locs = pnp_problem(args.views, rng)
    dt, (rot, loc) = timed(lambda: eng.pnp_refine(rowptr, xy, X, rots, locs, 5.0, args.pnp_iterations))
    print(json.dumps({'workload': 'C3 a19 batched nonlinear PnP', 'views': args.views, 'observations': int(rowptr[-1]),
                      'iterations': args.pnp_iterations, 'seconds': dt,
                      'observation_iterations_per_sec': int(rowptr[-1]) * args.pnp_iterations / dt,
                      'mean_abs_loc_move': float(np.mean(np.abs(loc - locs)))}))

    # f3: RANSAC hypotheses in parallel (host buffers; the reference loops over hypotheses and, inside, over points)
    n, H = 100000, 2000
    k = np.array([[800.0, 0, 320], [0, 800, 240], [0, 0, 1]])
    ang = 0.3
    rot0 = np.array([[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]])
    loc0 = np.array([[0.5], [-0.2], [0.1]])
    Xh = np.vstack([rng.uniform(-2, 2, (2, n)), rng.uniform(4, 9, (1, n)), np.ones((1, n))])
    pp = k @ np.hstack((rot0.T, rot0.T @ -loc0)) @ Xh
    key = pp / pp[2]
    bad = rng.random(n) < 0.3
    key[0:2, bad] += rng.normal(0, 40, (2, int(bad.sum())))
    samples = np.array([rng.choice(n, 6, replace=False) for _ in range(H)], dtype=np.int32)
    keyT, XT = np.ascontiguousarray(key.T), np.ascontiguousarray(Xh.T)
    dt, res = timed(lambda: eng.pnp_ransac(keyT, XT, k, samples, 1.0))
    print(json.dumps({'workload': 'f3 PnP RANSAC (6-point DLT), all hypotheses at once', 'points': n, 'hypotheses': H,
                      'seconds': dt, 'hypothesis_point_tests_per_sec': n * H / dt, 'inliers': res['count'],
                      'true_inliers': int((~bad).sum())}))
    left = rng.uniform(-1, 1, (n, 2))
    Ftrue = np.array([[0.0, -0.3, 0.2], [0.3, 0.0, -1.0], [-0.2, 1.0, 0.0]]) @ rot0
    ln = (Ftrue @ np.column_stack((left, np.ones(n))).T).T                 # epipolar lines in the right image
    xr = rng.uniform(-1, 1, n)
    yr = -(ln[:, 0] * xr + ln[:, 2]) / ln[:, 1]
    pairs = np.column_stack((left, xr, yr))
    pairs[bad, 3] += rng.normal(0, 0.2, int(bad.sum()))
    samples8 = np.array([rng.choice(n, 8, replace=False) for _ in range(H)], dtype=np.int32)
    dt, res = timed(lambda: eng.fund_ransac(pairs, samples8, 1e-6))
    print(json.dumps({'workload': 'f3 fundamental-matrix RANSAC (8-point), all hypotheses at once', 'pairs': n,
                      'hypotheses': H, 'seconds': dt, 'hypothesis_pair_tests_per_sec': n * H / dt,
                      'inliers': res['count'], 'true_inliers': int((~bad).sum())}))

    if not args.skip_ba:
        import torch
        for shape in ('ladybug', 'final'):
            prob, _ = synthetic.make_shape(shape)
            t = eng.to_device(prob)
            c0, p0 = t['cams'].clone(), t['pts'].clone()
            best, res = 1e30, None
            for _ in range(4):
                t['cams'].copy_(c0)
                t['pts'].copy_(p0)
                torch.cuda.synchronize()
                res = eng.ba_solve_device(t, lam=5.0, iterations=3)
                best = min(best, res.ms['total'])
            print(json.dumps({'workload': 'BA %s shape (reference mode, 3 LM iterations)' % shape, 'cams': prob.num_cams,
                              'pts': prob.num_pts, 'obs': prob.num_obs, 'device_ms': best,
                              'lm_iterations_per_sec': 3e3 / best, 'pcg_iterations': res.pcg_iterations,
                              'initial_cost': res.initial_cost, 'final_cost': res.final_cost,
                              'phase_ms': res.ms}))
            del t, c0, p0
            torch.cuda.empty_cache()
    eng.close()


if __name__ == '__main__':
    main()
