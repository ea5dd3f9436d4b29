"""Dense-Cholesky solver (sfmba_options.solver = DENSE_CHOLESKY): the explicit reduced camera system
S = A + lambda I - B D^-1 B^T of ba_processor.py:376-382, factored on the FP64 tensor cores.  Checked
against the golden fixtures of the unmodified reference, the oracle, and the PCG solver."""
import numpy as np
import pytest

from helpers import ba_case_from_npz, close, load_golden, max_rel, problem_from_case

pytestmark = pytest.mark.gpu

COST_RTOL = 1e-6
PARAM_TOL = 1e-5
KNOWN_COSTS = [2.069897342, 2.501760067, 4.348007169, 6.057589389]


def _dense():
    from sfm_b200 import _capi
    return _capi.SOLVER_DENSE_CHOLESKY


@pytest.mark.parametrize('call', [1, 4])
def test_upenn_ba_call_dense_matches_reference(engine, oracle, call):
    """The reference inverts the 7V x 7V reduced matrix directly; so does this mode (V = 3 and 6:
    one padded 128 x 128 tile)."""
    case = ba_case_from_npz(load_golden('upenn_ba_call%d.npz' % call))
    prob, _, _ = problem_from_case(case)
    res = engine.ba_solve_host(prob, lam=case['damping_factor'], iterations=case['iteration'], solver=_dense())
    rot = oracle.quaternion_to_rotation(prob.cams[:, 3:7])
    assert close(rot, case['out_rot'], PARAM_TOL), max_rel(rot, case['out_rot'])
    assert close(prob.cams[:, 0:3], case['out_loc'], PARAM_TOL)
    assert close(prob.pts, case['out_tri_pts'][0:3].T, PARAM_TOL)
    assert abs(res.final_cost - KNOWN_COSTS[call - 1]) <= COST_RTOL * KNOWN_COSTS[call - 1]
    assert res.pcg_iterations == 0 and res.flags == 0


@pytest.mark.parametrize('shape,iters,lam', [('tiny', 3, 5.0), ('small', 4, 1.0), ('ladybug', 3, 5.0)])
def test_dense_matches_oracle(engine, oracle, shape, iters, lam):
    """'small' (n = 7 V > 128) and 'ladybug' (343) exercise the panel solve and the trailing update."""
    from sfm_b200 import synthetic
    prob, _ = synthetic.make_shape(shape)
    b = prob.obs_uv.T.astype(np.float64)
    cams, pts, costs = oracle.ba_solve(prob.cams, prob.pts, prob.obs_cam, prob.obs_pt, b, lam, iters)
    res = engine.ba_solve_host(prob, lam=lam, iterations=iters, solver=_dense())
    assert close(prob.cams, cams, PARAM_TOL), max_rel(prob.cams, cams)
    assert close(prob.pts, pts, PARAM_TOL), max_rel(prob.pts, pts)
    assert abs(res.final_cost - costs[-1]) <= COST_RTOL * costs[-1]


def test_dense_agrees_with_pcg_on_many_cameras(engine):
    """700 cameras (n = 4900, 39 panels), every point seen by all of them (tracks longer than the
    32-observation tile of the assembly kernel), and a 300-camera banded problem: the direct solve and
    PCG run to 1e-13 give the same step."""
    from sfm_b200 import synthetic
    for make in (lambda: synthetic.make_problem(700, 40, 40 * 700, seed=3),
                 lambda: synthetic.make_problem(300, 20000, 100000, seed=5)):
        p1, _ = make()
        p2, _ = make()
        start = p1.cams.copy()
        r1 = engine.ba_solve_host(p1, lam=5.0, iterations=2, pcg_tol=1e-13)
        r2 = engine.ba_solve_host(p2, lam=5.0, iterations=2, solver=_dense())
        scale = np.max(np.abs(p1.cams - start))
        assert np.max(np.abs(p1.cams - p2.cams)) <= 1e-8 * max(scale, 1.0), np.max(np.abs(p1.cams - p2.cams))
        assert np.max(np.abs(p1.pts - p2.pts)) <= 1e-8 * max(1.0, np.max(np.abs(p1.pts)))
        assert abs(r1.final_cost - r2.final_cost) <= 1e-9 * r1.final_cost


def test_dense_adaptive_lm_and_unsorted_views(engine, oracle):
    """Adaptive damping with the direct solve; and a problem whose observations are not view-minor
    inside a point (the assembly kernel then writes the transposed block)."""
    from sfm_b200 import _capi, synthetic
    from sfm_b200.engine import BAProblem
    p, _ = synthetic.make_shape('small')
    res = engine.ba_solve_host(p, lam=5.0, iterations=6, mode=_capi.MODE_ADAPTIVE_LM, dvdc_sign_quirk=False,
                               solver=_dense())
    assert res.final_cost < res.initial_cost and res.iterations_accepted >= 1
    base, _ = synthetic.make_shape('small')
    perm = np.arange(base.num_obs)
    for j in range(base.num_pts):                          # reverse every point's observation list
        perm[base.rowptr[j]:base.rowptr[j + 1]] = perm[base.rowptr[j]:base.rowptr[j + 1]][::-1]
    q = BAProblem(cams=base.cams.copy(), pts=base.pts.copy(), obs_uv=np.ascontiguousarray(base.obs_uv[:, perm]),
                  obs_cam=np.ascontiguousarray(base.obs_cam[perm]), obs_pt=base.obs_pt, rowptr=base.rowptr)
    engine.ba_solve_host(base, lam=5.0, iterations=2, solver=_dense())
    engine.ba_solve_host(q, lam=5.0, iterations=2, solver=_dense())
    assert close(q.cams, base.cams, 1e-9) and close(q.pts, base.pts, 1e-9)
