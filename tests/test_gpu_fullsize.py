"""Full-size (Venice shape: 1 778 cameras / 993 923 points / 5 001 946 observations) checks of the CUDA
path through properties that do not need an oracle run at that size, plus a quarter-scale Venice
problem checked against the oracle's sparse solve, plus the degenerate inputs of the domain."""
import numpy as np
import pytest

from helpers import close, max_rel

pytestmark = pytest.mark.gpu

COST_RTOL = 1e-6
PARAM_TOL = 1e-5


@pytest.fixture(scope='module')
def venice():
    from sfm_b200 import synthetic
    prob, _ = synthetic.make_shape('venice')
    return prob


def _copy(p):
    from sfm_b200.engine import BAProblem
    return BAProblem(cams=p.cams.copy(), pts=p.pts.copy(), obs_uv=p.obs_uv, obs_cam=p.obs_cam, obs_pt=p.obs_pt,
                     rowptr=p.rowptr)


def test_venice_two_cost_paths_agree(engine, venice):
    """The index-free cost kernel (one thread per observation, obs_pt) and the tiled cost of the solve
    (camera-sorted runs inside point groups) are different code over the same 5 M observations."""
    direct = engine.ba_cost(venice)
    p = _copy(venice)
    res = engine.ba_solve_host(p, lam=5.0, iterations=1)
    assert abs(res.initial_cost - direct) <= 1e-11 * direct
    assert abs(res.final_cost - engine.ba_cost(p)) <= 1e-11 * res.final_cost


def test_venice_solve_is_bit_reproducible(engine, venice):
    """Every reduction has a fixed order: two solves of the same problem give identical bits."""
    p1, p2 = _copy(venice), _copy(venice)
    r1 = engine.ba_solve_host(p1, lam=5.0, iterations=2)
    r2 = engine.ba_solve_host(p2, lam=5.0, iterations=2)
    assert np.array_equal(p1.cams, p2.cams) and np.array_equal(p1.pts, p2.pts)
    assert r1.final_cost == r2.final_cost and r1.pcg_iterations == r2.pcg_iterations


def test_venice_reference_mode_step_solves_the_normal_equations(engine, venice):
    """One reference-mode iteration from two PCG tolerances: the steps agree to the looser tolerance,
    i.e. the reduced system really is solved (the step is not an artefact of the stopping rule)."""
    from sfm_b200 import _capi
    p1, p2, p3 = _copy(venice), _copy(venice), _copy(venice)
    engine.ba_solve_host(p1, lam=5.0, iterations=1, pcg_tol=1e-13)
    engine.ba_solve_host(p2, lam=5.0, iterations=1, pcg_tol=1e-8, matvec_precision=_capi.MATVEC_F64)
    step1 = p1.cams - venice.cams
    scale = np.max(np.abs(step1))
    pscale = max(1.0, np.max(np.abs(p1.pts - venice.pts)))
    assert np.max(np.abs(p1.cams - p2.cams)) <= 1e-6 * scale
    assert np.max(np.abs(p1.pts - p2.pts)) <= 1e-6 * pscale
    # the default precision (FP32 Jacobian arithmetic and matvec, FP64 sums) on top of the default tolerance: the step
    # carries the 1e-7 of its arithmetic, an order inside the 1e-5 the parameters are held to
    engine.ba_solve_host(p3, lam=5.0, iterations=1, pcg_tol=1e-8)
    assert np.max(np.abs(p1.cams - p3.cams)) <= 1e-5 * scale
    assert np.max(np.abs(p1.pts - p3.pts)) <= 1e-5 * pscale


def test_venice_adaptive_lm_without_quirk_decreases_cost(engine, venice):
    """With the exact dv/dC Jacobian and adaptive damping every accepted step lowers the cost."""
    from sfm_b200 import _capi
    p = _copy(venice)
    res = engine.ba_solve_host(p, lam=5.0, iterations=8, mode=_capi.MODE_ADAPTIVE_LM, dvdc_sign_quirk=False)
    assert res.final_cost < res.initial_cost
    assert res.iterations_accepted >= 1
    accepted = [c for c in res.costs if c > 0]
    assert all(b <= a for a, b in zip(accepted, accepted[1:]))          # the cost history never rises
    assert abs(res.final_cost - engine.ba_cost(p)) <= 1e-11 * res.final_cost


def test_quarter_venice_matches_oracle(engine, oracle):
    """250 k points / 1.25 M observations, one reference-mode iteration against the oracle's sparse solve."""
    from sfm_b200 import synthetic
    prob, _ = synthetic.make_shape('venice', scale=0.25)
    b = prob.obs_uv.T.astype(np.float64)
    cams, pts, costs = oracle.ba_solve(prob.cams, prob.pts, prob.obs_cam, prob.obs_pt, b, 5.0, 1, sparse=True)
    res = engine.ba_solve_host(prob, lam=5.0, iterations=1)
    assert close(prob.cams, cams, PARAM_TOL), max_rel(prob.cams, cams)
    assert close(prob.pts, pts, PARAM_TOL), max_rel(prob.pts, pts)
    assert abs(res.initial_cost - costs[0]) <= COST_RTOL * costs[0]
    assert abs(res.final_cost - costs[-1]) <= COST_RTOL * costs[-1]


def test_degenerate_inputs(engine, oracle):
    """No points at all, one camera, and a single two-view point."""
    from sfm_b200 import synthetic
    from sfm_b200.engine import BAProblem
    base, _ = synthetic.make_shape('tiny')
    # no points, no observations: nothing moves, cost 0
    p = BAProblem(cams=base.cams.copy(), pts=np.zeros((0, 3)), obs_uv=np.zeros((2, 0), np.float32),
                  obs_cam=np.zeros(0, np.int32), obs_pt=np.zeros(0, np.int32), rowptr=np.zeros(1, np.int32))
    res = engine.ba_solve_host(p, lam=5.0, iterations=2)
    assert res.final_cost == 0.0 and np.array_equal(p.cams[:, 0:3], base.cams[:, 0:3])
    # one two-view point
    sel = slice(base.rowptr[0], base.rowptr[1])
    keep = base.obs_cam[sel][:2]
    p = BAProblem(cams=base.cams.copy(), pts=base.pts[0:1].copy(), obs_uv=base.obs_uv[:, sel][:, :2].copy(),
                  obs_cam=keep.copy(), obs_pt=np.zeros(2, np.int32), rowptr=np.array([0, 2], np.int32))
    b = p.obs_uv.T.astype(np.float64)
    cams, pts, costs = oracle.ba_solve(p.cams, p.pts, p.obs_cam, p.obs_pt, b, 5.0, 2)
    res = engine.ba_solve_host(p, lam=5.0, iterations=2)
    assert close(p.cams, cams, PARAM_TOL) and close(p.pts, pts, PARAM_TOL)
    assert abs(res.final_cost - costs[-1]) <= COST_RTOL * max(costs[-1], 1e-30)


def test_venice_host_and_device_paths_are_bit_identical(engine, venice):
    """The host-buffer solve uploads Venice in four ranges of whole groups and builds the records / runs the first
    linearisation range by range behind the copies; the device-resident solve does it in one launch each.  Same
    kernels over the same groups, fixed-order sums: identical bits, in both LM modes."""
    from sfm_b200 import _capi
    for kw in (dict(lam=5.0, iterations=2), dict(lam=1e-3, iterations=3, mode=_capi.MODE_ADAPTIVE_LM,
                                                  dvdc_sign_quirk=False)):
        ph = _copy(venice)
        rh = engine.ba_solve_host(ph, **kw)
        t = engine.to_device(venice)
        rd = engine.ba_solve_device(t, **kw)
        assert np.array_equal(ph.cams, t['cams'].cpu().numpy())
        assert np.array_equal(ph.pts, t['pts'].cpu().numpy())
        assert rh.final_cost == rd.final_cost and rh.initial_cost == rd.initial_cost
        assert rh.costs == rd.costs


def test_final_shape_properties(engine):
    """The largest configuration (Final shape: 13 682 cameras / 4 456 117 points / 28 987 644 observations) on one GPU,
    through the same size-independent properties: the two cost paths agree, the host-buffer (ranged upload) and the
    device-resident solves give identical bits, and the reference-mode step solves the normal equations."""
    from sfm_b200 import synthetic
    final, _ = synthetic.make_shape('final')
    direct = engine.ba_cost(final)
    ph = _copy(final)
    rh = engine.ba_solve_host(ph, lam=5.0, iterations=2)
    assert abs(rh.initial_cost - direct) <= 1e-11 * direct
    assert abs(rh.final_cost - engine.ba_cost(ph)) <= 1e-11 * rh.final_cost
    t = engine.to_device(final)
    rd = engine.ba_solve_device(t, lam=5.0, iterations=2)
    assert np.array_equal(ph.cams, t['cams'].cpu().numpy())
    assert np.array_equal(ph.pts, t['pts'].cpu().numpy())
    assert rh.costs == rd.costs and rh.pcg_iterations == rd.pcg_iterations
    from sfm_b200 import _capi
    p1, p2, p3 = _copy(final), _copy(final), _copy(final)
    engine.ba_solve_host(p1, lam=5.0, iterations=1, pcg_tol=1e-13)
    engine.ba_solve_host(p2, lam=5.0, iterations=1, pcg_tol=1e-8, matvec_precision=_capi.MATVEC_F64)
    engine.ba_solve_host(p3, lam=5.0, iterations=1, pcg_tol=1e-8)         # default precision: FP32 arithmetic, FP64 sums
    scale = np.max(np.abs(p1.cams - final.cams))
    pscale = max(1.0, np.max(np.abs(p1.pts - final.pts)))
    assert np.max(np.abs(p1.cams - p2.cams)) <= 1e-6 * scale
    assert np.max(np.abs(p1.pts - p2.pts)) <= 1e-6 * pscale
    assert np.max(np.abs(p1.cams - p3.cams)) <= 1e-5 * scale
    assert np.max(np.abs(p1.pts - p3.pts)) <= 1e-5 * pscale
