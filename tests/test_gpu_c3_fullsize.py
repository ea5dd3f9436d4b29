"""BASELINE.json config C3 at its stated size -- 1 M points (a18) / 10 k views (a19) in one batched call each -- checked
against the oracle on a fixed random sample of the batch (>= 1 000 points, >= 100 views) with the reference's own
iteration counts and damping (triangulation_processor.py:12 / 160-234: lambda 0.5, 100 iterations;
campose_processor.py:308-459 as the pipeline calls it: lambda 5, 300 iterations)."""
import numpy as np
import pytest

from helpers import close, max_rel

pytestmark = pytest.mark.gpu

TOL = 1e-8


def test_c3_triangulation_one_million_points(engine, oracle):
    from sfm_b200 import synthetic
    M = 1000000
    projs, rowptr, obs_view, uv, X0 = synthetic.tri_problem(M, np.random.default_rng(3))
    out = engine.triangulate_refine(projs, rowptr, obs_view, uv, X0, 0.5, 100)
    assert out.shape == (M, 3) and np.all(np.isfinite(out))
    sample = np.random.default_rng(11).choice(M, 1000, replace=False)
    worst = 0.0
    for j in sample:
        o = slice(rowptr[j], rowptr[j + 1])
        init = np.append(X0[j], 1.0).reshape(4, 1)
        pairs = [np.array([[uv[q, 0]], [uv[q, 1]], [1.0]]) for q in range(o.start, o.stop)]
        ref = oracle.nonlinear_triangulate(init, projs[obs_view[o]], pairs, 0.5, 100)
        worst = max(worst, max_rel(out[j], ref[0:3, 0]))
        assert close(out[j], ref[0:3, 0], TOL), (j, out[j], ref[0:3, 0])
    print('C3 a18: 1000 sampled points of 1 M, worst relative difference to the oracle %.2e' % worst)
    # every track length 2..6 is in the sample
    assert set(np.diff(rowptr)[sample].tolist()) == {2, 3, 4, 5, 6}


def test_c3_pnp_ten_thousand_views(engine, oracle):
    from sfm_b200 import synthetic
    P = 10000
    rowptr, xy, X, rots, locs = synthetic.pnp_problem(P, np.random.default_rng(4))
    rot, loc = engine.pnp_refine(rowptr, xy, X, rots, locs, 5.0, 300)
    assert rot.shape == (P, 3, 3) and np.all(np.isfinite(rot)) and np.all(np.isfinite(loc))
    sample = np.random.default_rng(12).choice(P, 100, replace=False)
    worst = 0.0
    for v in sample:
        o = slice(rowptr[v], rowptr[v + 1])
        n = o.stop - o.start
        key = np.vstack((xy[o].T, np.ones(n)))
        tri = np.vstack((X[o].T, np.ones(n)))
        r_ref, l_ref = oracle.nonlinear_pnp(key, tri, np.eye(3), rots[v], locs[v].reshape(3, 1), 5.0, 300)
        worst = max(worst, max_rel(rot[v], r_ref), max_rel(loc[v], l_ref.reshape(3)))
        assert close(rot[v], r_ref, TOL) and close(loc[v], l_ref.reshape(3), TOL), v
    print('C3 a19: 100 sampled views of 10 k, worst relative difference to the oracle %.2e' % worst)
