"""Huber loss (sfmba_options.huber_delta; SURVEY.md section 8f row 4 -- not in the reference): the GPU's re-weighted
normal equations against the oracle's restatement, and the adaptive LM with outliers."""
import numpy as np
import pytest

from helpers import close, max_rel

pytestmark = pytest.mark.gpu


def _with_outliers(shape, frac, seed):
    from sfm_b200 import synthetic
    prob, truth = synthetic.make_shape(shape)
    rng = np.random.default_rng(seed)
    bad = rng.random(prob.num_obs) < frac
    prob.obs_uv[:, bad] += rng.normal(0, 0.05, (2, int(bad.sum()))).astype(np.float32)
    return prob, truth, bad


@pytest.mark.parametrize('shape,quirk', [('tiny', True), ('small', False), ('small', True)])
def test_huber_step_matches_oracle(engine, oracle, shape, quirk):
    """Reference-style fixed-lambda steps with the Huber weights: cameras, points and sum rho agree with the oracle."""
    prob, _, bad = _with_outliers(shape, 0.1, 21)
    delta = 0.004
    b = prob.obs_uv.T.astype(np.float64)
    cams, pts, costs = oracle.ba_solve(prob.cams, prob.pts, prob.obs_cam, prob.obs_pt, b, 5.0, 3, dvdc_quirk=quirk,
                                       huber_delta=delta)
    plain = oracle.ba_cost(prob.cams, prob.pts, prob.obs_cam, prob.obs_pt, b)
    res = engine.ba_solve_host(prob, lam=5.0, iterations=3, dvdc_sign_quirk=quirk, huber_delta=delta, pcg_tol=1e-13)
    assert close(prob.cams, cams, 1e-8), max_rel(prob.cams, cams)
    assert close(prob.pts, pts, 1e-8), max_rel(prob.pts, pts)
    assert abs(res.initial_cost - costs[0]) <= 1e-10 * costs[0]
    assert abs(res.final_cost - costs[-1]) <= 1e-9 * costs[-1]
    assert costs[0] < 0.8 * plain                      # the weights are active on this problem
    assert [abs(a - c) <= 1e-9 * c for a, c in zip(res.costs, costs)]


def test_huber_off_is_bit_identical(engine):
    from sfm_b200 import synthetic
    p1, _ = synthetic.make_shape('small')
    p2, _ = synthetic.make_shape('small')
    engine.ba_solve_host(p1, lam=5.0, iterations=3)
    engine.ba_solve_host(p2, lam=5.0, iterations=3, huber_delta=0.0)
    assert np.array_equal(p1.cams, p2.cams) and np.array_equal(p1.pts, p2.pts)
    with pytest.raises(Exception):
        engine.ba_solve_host(p2, lam=5.0, iterations=1, huber_delta=-1.0)


def test_adaptive_lm_with_huber_rejects_outliers(engine, oracle):
    """Adaptive LM, exact Jacobian: with 10 % gross outliers the Huber solve ends closer to the true cameras than the
    plain least-squares solve, its cost history is monotone, and huge delta reproduces plain least squares."""
    from sfm_b200 import _capi
    prob_h, truth, bad = _with_outliers('small', 0.1, 5)
    prob_l, _, _ = _with_outliers('small', 0.1, 5)
    prob_big, _, _ = _with_outliers('small', 0.1, 5)
    # (FP64 matvec in all three: the Huber solves always run it, and 25 adaptive iterations would otherwise compare the
    # FP32 matvec of the plain solve with the FP64 one of the huge-delta solve instead of the two losses)
    kw = dict(lam=1e-3, iterations=25, mode=_capi.MODE_ADAPTIVE_LM, dvdc_sign_quirk=False,
              matvec_precision=_capi.MATVEC_F64)
    rh = engine.ba_solve_host(prob_h, huber_delta=0.002, **kw)
    rl = engine.ba_solve_host(prob_l, **kw)
    rb = engine.ba_solve_host(prob_big, huber_delta=1e6, **kw)
    assert rh.final_cost < rh.initial_cost and rl.final_cost < rl.initial_cost
    assert close(prob_big.cams, prob_l.cams, 1e-9) and abs(rb.final_cost - rl.final_cost) <= 1e-9 * rl.final_cost
    # residuals of the clean observations at the two solutions
    b = prob_h.obs_uv.T.astype(np.float64)

    def clean_rms(p):
        r, _, _ = oracle.residual_jacobian(p.cams, p.pts, p.obs_cam, p.obs_pt, b, want_jac=False)
        return float(np.sqrt(np.mean(np.sum(r[~bad] ** 2, axis=1))))
    assert clean_rms(prob_h) < 0.7 * clean_rms(prob_l), (clean_rms(prob_h), clean_rms(prob_l))


def test_adaptive_lm_gain_ratio_and_damping_match_a_numpy_lm(engine, oracle):
    """The device-side LM bookkeeping (k_lm_model_decrease, k_lm_decide): predicted decrease delta . J^T r + lambda |delta|^2
    over cameras AND points, gain ratio, Nielsen's damping update, accept / reject -- replayed in NumPy on the oracle's
    normal equations with the exact reduced solve.  Same accepted steps, same final damping, same cost."""
    from sfm_b200 import _capi, synthetic
    prob, _ = synthetic.make_shape('small')
    b = prob.obs_uv.T.astype(np.float64)
    oc, op = prob.obs_cam.astype(np.int64), prob.obs_pt.astype(np.int64)
    cams, pts = prob.cams.copy(), prob.pts.copy()
    V, N = cams.shape[0], pts.shape[0]
    lam, nu, accepted, iters = 1e-2, 2.0, 0, 6
    for _ in range(iters):
        ne = oracle.ba_normal_equations(cams, pts, oc, op, b, lam, dvdc_quirk=False)
        S, g = oracle._reduced_system_dense(ne, oc, op, V)
        dp = np.linalg.solve(S, g).reshape(V, 7)
        bt_dp = oracle._segment_sum(np.einsum('mij,mi->mj', ne['W'], dp[oc]), op, N)
        dx = np.einsum('nij,nj->ni', ne['Dinv'], ne['e_x'] - bt_dp)
        pred = float(np.sum(dp * ne['e_p']) + np.sum(dx * ne['e_x']) + lam * (np.sum(dp * dp) + np.sum(dx * dx)))
        c2 = cams + dp
        c2[:, 3:7] /= np.sqrt(np.sum(np.square(c2[:, 3:7]), axis=1, keepdims=True))
        p2 = pts + dx
        new = oracle.ba_cost(c2, p2, oc, op, b)
        if new < ne['cost']:
            rho = (ne['cost'] - new) / pred if pred > 0 else 1.0
            lam *= max(1.0 / 3.0, 1.0 - (2.0 * rho - 1.0) ** 3)
            nu = 2.0
            cams, pts = c2, p2
            accepted += 1
        else:
            lam *= nu
            nu *= 2.0
    res = engine.ba_solve_host(prob, lam=1e-2, iterations=iters, mode=_capi.MODE_ADAPTIVE_LM, dvdc_sign_quirk=False,
                               pcg_tol=1e-13)
    assert res.iterations_accepted == accepted
    assert abs(res.final_lambda - lam) <= 1e-6 * lam, (res.final_lambda, lam)
    assert close(prob.cams, cams, 1e-7) and close(prob.pts, pts, 1e-7)
    assert abs(res.final_cost - oracle.ba_cost(cams, pts, oc, op, b)) <= 1e-8 * res.final_cost
