"""Pins the CPU oracle (oracle/sfm_oracle.py) against fixtures recorded from the UNMODIFIED reference
(tests/golden/make_golden.py) and against the known answers of SURVEY.md section 8(c).  CPU only."""
import numpy as np
import pytest

from helpers import ba_case_from_npz, load_golden

KNOWN_COSTS = [2.069897342, 2.501760067, 4.348007169, 6.057589389]


def _oracle_inputs(oracle, case):
    """Packs one recorded BA call the way the oracle wants it (no product code involved)."""
    V, N = case['V'], case['N']
    obs_pt, obs_cam, obs_key = oracle.build_visibility(case['diag'], N)
    cams = np.zeros((V, 7))
    cams[:, 0:3] = case['loc']
    cams[:, 3:7] = oracle.rotation_to_quaternion(case['rot'])
    px = np.zeros((obs_pt.shape[0], 2))
    for v in range(V):
        sel = obs_cam == v
        # cv2.KeyPoint.pt is a float32 pair
        px[sel] = case['keys'][v][obs_key[sel]].astype(np.float32).astype(np.float64)
    b = oracle.normalise_observations(px, obs_cam, case['k'])
    pts = case['tri_pts'][0:3].T.copy()
    return cams, pts, obs_cam, obs_pt, obs_key, b


@pytest.mark.parametrize('call', [1, 2, 3, 4])
def test_oracle_reproduces_upenn_ba_call(oracle, call):
    case = ba_case_from_npz(load_golden('upenn_ba_call%d.npz' % call))
    cams, pts, obs_cam, obs_pt, obs_key, b = _oracle_inputs(oracle, case)
    # bit-exact loop-order index (ba_processor.py:304-310 + key_tracker.py:198-204)
    assert np.array_equal(np.stack([obs_pt, obs_cam, obs_key], 1), case['vis'])
    # first linearisation, block by block
    r, jp, jx = oracle.residual_jacobian(cams, pts, obs_cam, obs_pt, b)
    assert np.max(np.abs(b - case['b0'])) < 1e-14
    assert np.max(np.abs(r - (case['b0'] - case['f0']))) < 1e-12
    assert np.max(np.abs(jp - case['jp0'])) < 1e-9 * max(1.0, np.max(np.abs(case['jp0'])))
    assert np.max(np.abs(jx - case['jx0'])) < 1e-10
    # whole call
    oc, op, costs = oracle.ba_solve(cams, pts, obs_cam, obs_pt, b, case['damping_factor'], case['iteration'])
    rot = oracle.quaternion_to_rotation(oc[:, 3:7])
    assert np.max(np.abs(rot - case['out_rot'])) < 1e-10
    assert np.max(np.abs(oc[:, 0:3] - case['out_loc'])) < 1e-10
    ref = case['out_tri_pts'][0:3].T
    assert np.max(np.abs(op - ref) / np.maximum(1.0, np.abs(ref))) < 1e-10
    assert abs(costs[-1] - KNOWN_COSTS[call - 1]) < 2e-9


def test_oracle_reproduces_literal_reference_on_random_problems(oracle):
    z = load_golden('synth_ba_small.npz')
    assert int(z['num_cases']) >= 2
    for c in range(int(z['num_cases'])):
        case = ba_case_from_npz(z, 'c%d_' % c)
        cams, pts, obs_cam, obs_pt, obs_key, b = _oracle_inputs(oracle, case)
        assert np.array_equal(np.stack([obs_pt, obs_cam, obs_key], 1), case['vis'])
        for sparse in (False, True):
            oc, op, _ = oracle.ba_solve(cams, pts, obs_cam, obs_pt, b, case['damping_factor'], case['iteration'],
                                        sparse=sparse)
            assert np.max(np.abs(oracle.quaternion_to_rotation(oc[:, 3:7]) - case['out_rot'])) < 1e-9
            assert np.max(np.abs(oc[:, 0:3] - case['out_loc'])) < 1e-9
            assert np.max(np.abs(op - case['out_tri_pts'][0:3].T)) < 1e-8


def test_pipeline_end_state_matches_stored_results():
    """The container run of the reference reproduces test_dataset/upenn/results/*.pkl."""
    z = load_golden('upenn_final.npz')
    assert np.max(np.abs(z['loc'] - z['stored_loc'])) < 1e-9
    assert np.max(np.abs(z['rot'] - z['stored_rot'])) < 1e-9
    ref = z['stored_tri_pts']
    assert np.max(np.abs(z['tri_pts'] - ref) / np.maximum(1.0, np.abs(ref))) < 1e-8
    # and the last BA call's recorded output is that end state
    case = ba_case_from_npz(load_golden('upenn_ba_call4.npz'))
    assert np.array_equal(case['out_rot'], z['rot'])
    assert np.array_equal(case['out_tri_pts'], z['tri_pts'])


def test_is_visible_quirk_against_literal(oracle):
    rng = np.random.default_rng(7)
    N = 40
    rows = []
    for v in range(5):
        row = rng.integers(-1, N, size=60)
        row[rng.random(60) < 0.5] = -1
        rows.append(row)
    rows[0][0] = 3          # only match of point 3 in view 0 is key 0 -> invisible (np.any quirk)
    rows[0][rows[0] == 3] = -1
    rows[0][0] = 3
    rows[1][0] = 5          # key 0 and a second key -> visible through key 0
    rows[1][7] = 5
    obs_pt, obs_cam, obs_key = oracle.build_visibility(rows, N)
    lit = [(t, v, oracle.is_visible_literal(rows[v], t)) for t in range(N) for v in range(5)]
    lit = [(t, v, k) for t, v, k in lit if k != -1]
    assert [tuple(x) for x in np.stack([obs_pt, obs_cam, obs_key], 1).tolist()] == lit
    assert (3, 0) not in {(t, v) for t, v, _ in lit}
    assert (5, 1, 0) in lit


def test_oracle_triangulation_known_answers(oracle):
    z = load_golden('tri_inline.npz')
    projs, pairs = z['projs'], z['pairs']
    out2 = oracle.nonlinear_triangulate(z['linear'], [projs[0], projs[1]], [pairs[0], pairs[1]], 0.5, 300)
    assert np.max(np.abs(out2 - z['out2'])) < 1e-11
    # triangulation_processor.py:415-473 prints [-0.03470014 -0.0059831 2.02198114]
    assert np.allclose(out2[0:3, 0], [-0.03470014, -0.0059831, 2.02198114], atol=5e-9)
    jac, _ = oracle.triangulation_jacobian(z['linear'][0:3].T, projs)
    assert np.max(np.abs(jac[0] - z['jac3'])) < 1e-9 * np.max(np.abs(z['jac3']))
    calls = load_golden('upenn_tri_calls.npz')
    for c in range(int(calls['num_calls'])):
        g = lambda k: calls['c%d_%s' % (c, k)]  # noqa: E731
        out = oracle.nonlinear_triangulate(g('init'), list(g('projs')), list(g('pairs')),
                                           float(g('damping_factor')), int(g('iteration')))
        ref = g('out')
        assert np.max(np.abs(out - ref) / np.maximum(1.0, np.abs(ref))) < 1e-9


def test_oracle_linear_triangulation_reproduces_reference(oracle):
    """f2: every linear_triangulate result of the upenn pipeline (the `init` argument the reference passed on
    to nonlinear_triangulate) and the inline known answer (triangulation_processor.py:415-473)."""
    z = load_golden('upenn_tri_calls.npz')
    for c in range(5):
        lin = oracle.linear_triangulate(list(z['c%d_projs' % c]), list(z['c%d_pairs' % c]))
        ref = z['c%d_init' % c]
        assert lin.shape == ref.shape
        assert np.max(np.abs(lin - ref) / np.maximum(1.0, np.abs(ref))) < 1e-12
        front = oracle.cheirality_indices(z['c%d_projs' % c][0], z['c%d_projs' % c][1], ref)
        d1 = (z['c%d_projs' % c][0] @ ref)[2]
        d2 = (z['c%d_projs' % c][1] @ ref)[2]
        assert front == [i for i in range(ref.shape[1]) if d1[i] > 0 and d2[i] > 0]      # campose_processor.py:181-189
    t = load_golden('tri_inline.npz')
    lin = oracle.linear_triangulate(list(t['projs'][:2]), list(t['pairs'][:2]))
    assert np.max(np.abs(lin - t['linear'])) < 1e-12
    assert np.allclose(lin[0:3, 0], [-0.0347069, -0.00593501, 2.02198105], atol=5e-9)


def test_oracle_pnp_known_answers(oracle):
    z = load_golden('opencv_pnp.npz')
    rot, loc = oracle.nonlinear_pnp(z['key_2d'], z['tri_3d'], z['k'], z['init_rot'], z['init_loc'],
                                    float(z['damping_factor']), int(z['iteration']))
    assert np.max(np.abs(rot - z['out_rot'])) < 1e-9
    assert np.max(np.abs(loc - z['out_loc'])) < 1e-9
    # campose_processor.py:1073-1090: within 1 degree / 0.1 of the stored OpenCV pose
    assert np.sqrt(np.sum((loc - z['loc_truth']) ** 2)) < 0.1
    calls = load_golden('upenn_pnp_calls.npz')
    for c in range(int(calls['num_calls'])):
        g = lambda k: calls['c%d_%s' % (c, k)]  # noqa: E731
        rot, loc = oracle.nonlinear_pnp(g('key_2d'), g('tri_3d'), g('k'), g('init_rot'), g('init_loc'),
                                        float(g('damping_factor')), int(g('iteration')))
        assert np.max(np.abs(rot - g('out_rot'))) < 1e-9
        assert np.max(np.abs(loc - g('out_loc'))) < 1e-8


def test_jacobian_quirk_vs_finite_differences(oracle):
    """Only the three dv/dC entries of the reference Jacobian are wrong."""
    rng = np.random.default_rng(3)
    q = rng.normal(size=4) * 0.2
    q[0] = 1.0
    q /= np.linalg.norm(q)
    cam = np.concatenate([rng.normal(size=3), q])
    X = np.array([0.3, -0.2, 6.0]) + cam[0:3]
    b = np.zeros((1, 2))
    idx = np.zeros(1, dtype=np.int64)

    def f(c, x):
        r, _, _ = oracle.residual_jacobian(c[None], x[None], idx, idx, b, want_jac=False)
        return -r[0]

    _, jp_fix, jx = oracle.residual_jacobian(cam[None], X[None], idx, idx, b, dvdc_quirk=False)
    _, jp_q, _ = oracle.residual_jacobian(cam[None], X[None], idx, idx, b, dvdc_quirk=True)
    h = 1e-6
    # the reference's quaternion Jacobian is that of the (non-normalised) polynomial q -> R(q)
    num = np.stack([(f(cam + h * e, X) - f(cam - h * e, X)) / (2 * h) for e in np.eye(7)], axis=1)
    numx = np.stack([(f(cam, X + h * e) - f(cam, X - h * e)) / (2 * h) for e in np.eye(3)], axis=1)
    assert np.max(np.abs(jx[0] - numx)) < 1e-7
    assert np.max(np.abs(jp_fix[0] - num)) < 1e-6
    diff = np.abs(jp_q[0] - jp_fix[0]) > 1e-12
    assert not diff[0].any() and not diff[1, 3:].any() and diff[1, 0:3].any()


# ---- f3: RANSAC initialisers ---------------------------------------------------------------------------
@pytest.mark.parametrize('tag', ['clean', 'outliers'])
def test_oracle_pnp_ransac_matches_reference(oracle, tag):
    """oracle.pnp_ransac on the sample table the reference drew == linear_estimate_cam_pose_pnp's own result
    (campose_processor.py:485-633), including the LAPACK-sign dependent (R, -C) hypotheses."""
    g = load_golden('ransac.npz')
    r = oracle.pnp_ransac(g['pnp_%s_key' % tag], g['pnp_tri'], g['pnp_k'], g['pnp_%s_samples' % tag],
                          float(g['pnp_%s_thr' % tag]))
    assert r['inliers'] == list(g['pnp_%s_inliers' % tag])
    assert np.max(np.abs(r['rot'] - g['pnp_%s_rot' % tag])) < 1e-12
    assert np.max(np.abs(r['loc'] - g['pnp_%s_loc' % tag])) < 1e-9
    frac = float(np.mean(r['signs'] > 0))
    assert 0.2 < frac < 0.8            # both signs occur: the quirk is real on this host's LAPACK


@pytest.mark.parametrize('tag', ['clean', 'outliers'])
def test_oracle_fund_ransac_matches_reference(oracle, tag):
    g = load_golden('ransac.npz')
    pairs, tl, tr = oracle.fund_normalize(g['fund_%s_left' % tag], g['fund_%s_right' % tag])
    r = oracle.fund_ransac(pairs, g['fund_%s_samples' % tag], float(g['fund_%s_thr' % tag]))
    assert r['inliers'] == list(g['fund_%s_inliers' % tag])
    f = oracle.fund_denormalize(r['F'], tl, tr)
    assert np.max(np.abs(f - g['fund_%s_F' % tag])) < 1e-10 * np.max(np.abs(f))


def test_oracle_fund_eight_known_answer(oracle):
    """epipolar_processor.py:274-299: eight pairs, no RANSAC."""
    g = load_golden('ransac.npz')
    d8 = g['fund_eight_data']
    pairs, tl, tr = oracle.fund_normalize(d8[:, 0:2], d8[:, 2:4])
    f = oracle.fund_denormalize(oracle.fund_eight(pairs), tl, tr)
    assert np.max(np.abs(f - g['fund_eight_F'])) < 1e-9 * np.max(np.abs(f))
