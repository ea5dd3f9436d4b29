"""The BENCHMARKED bundle-adjustment configurations against full-size oracle goldens (tests/golden/bal_*.npz, made by
tests/golden/make_golden_bal.py from oracle/sfm_oracle.py's exact reduced solve; ba_processor.py:297-406).  The GPU
solves run with the DEFAULT options (PCG to sfmba_options.pcg_tol = 1e-8, block-Jacobi on A + lambda I), through the
host-buffer C-ABI entry point, so the default tolerance is checked on the answer, not on the residual:
final cost <= 1e-6 relative, cameras and points <= 1e-5 (BASELINE.json north_star)."""
import numpy as np
import pytest

from helpers import load_golden, max_rel

pytestmark = pytest.mark.gpu

COST_RTOL = 1e-6
PARAM_TOL = 1e-5


def input_checksums(prob):
    return np.array([prob.cams.sum(), prob.pts.sum(), float(prob.obs_uv.astype(np.float64).sum()),
                     float(prob.obs_cam.astype(np.int64).sum()), float(prob.rowptr.astype(np.int64).sum())])


def _run_case(engine, name, **kw):
    from sfm_b200 import synthetic
    z = load_golden(name + '.npz')
    prob, _ = synthetic.make_shape(str(z['shape']), scale=float(z['scale']))
    assert [prob.num_cams, prob.num_pts, prob.num_obs] == z['sizes'].tolist()
    # the generator must reproduce the problem the golden was made from, to the bit
    assert np.array_equal(input_checksums(prob), z['input_checksums']), 'synthetic generator drifted from the golden'
    res = engine.ba_solve_host(prob, lam=float(z['lam']), iterations=int(z['iterations']),
                               dvdc_sign_quirk=bool(z['dvdc_quirk']), **kw)
    costs = z['costs']
    errs = dict(cams=max_rel(prob.cams, z['cams']), pts=max_rel(prob.pts[z['pts_idx']], z['pts_sample']),
                cost=abs(res.final_cost - costs[-1]) / costs[-1],
                pts_sum=max_rel(prob.pts.sum(axis=0), z['pts_sum']),
                pts_sumsq=max_rel((prob.pts * prob.pts).sum(axis=0), z['pts_sumsq']))
    print(name, errs, 'pcg iterations', res.pcg_iterations, 'flags', res.flags)
    assert res.flags == 0
    for k in range(int(z['iterations'])):
        assert abs(res.costs[k] - costs[k]) <= COST_RTOL * costs[k], (k, res.costs[k], costs[k])
    assert errs['cost'] <= COST_RTOL, errs
    assert errs['cams'] <= PARAM_TOL, errs
    assert errs['pts'] <= PARAM_TOL, errs
    # checksums over ALL points (every point enters, not only the stored sample)
    assert errs['pts_sum'] <= PARAM_TOL and errs['pts_sumsq'] <= PARAM_TOL, errs
    return res, errs


def test_venice_headline_config_matches_oracle(engine):
    """BASELINE.json C4 = bench.py's workload: full Venice, lambda 5, 3 reference-mode iterations, default options."""
    res, errs = _run_case(engine, 'bal_venice_ref')
    assert res.iterations_done == 3


def test_ladybug_config_matches_oracle(engine):
    """BASELINE.json C2."""
    _run_case(engine, 'bal_ladybug_ref')


def test_weak_damping_default_tolerance_matches_oracle(engine):
    """lambda = 1e-2 with the exact dv/dC Jacobian on quarter-Venice: the reduced system is far less diagonally
    dominant than at lambda = 5 (block-Jacobi PCG needs ~20 iterations instead of 3), which is where a loose default
    tolerance would show in the answer."""
    res, errs = _run_case(engine, 'bal_venice_weak')
    assert res.pcg_iterations > 2 * 6         # really the slow regime
