"""GPU parity of the batched refiners (a18, a19) and of the class-level drop-in API, against the
fixtures recorded from the unmodified reference and against the oracle."""
import numpy as np
import pytest

from helpers import (FakeTracker, FakeView, FakeViewProcessor, ba_case_from_npz, close, load_golden, max_rel)

pytestmark = pytest.mark.gpu
TOL = 1e-5


def test_triangulation_inline_known_answer(engine):
    """triangulation_processor.py:415-473: linear DLT then 300 iterations at lambda 0.5."""
    import sfm_b200
    z = load_golden('tri_inline.npz')
    tp = sfm_b200.TriangulationProcessor(engine=engine)
    pairs = [z['pairs'][0], z['pairs'][1]]
    lin = tp.linear_triangulate([z['projs'][0], z['projs'][1]], pairs)
    assert np.allclose(lin, z['linear'], atol=1e-9)
    out = tp.triangulate([z['projs'][0], z['projs'][1]], pairs, 0.5, 300)
    assert close(out, z['out2'], 1e-9), max_rel(out, z['out2'])
    assert np.allclose(out[0:3, 0], [-0.03470014, -0.0059831, 2.02198114], atol=1e-7)
    out3 = tp.nonlinear_triangulate(z['linear'], list(z['projs']), list(z['pairs']), 0.5, 300)
    assert close(out3, z['out3'], 1e-9)


def test_triangulation_pipeline_calls(engine):
    import sfm_b200
    z = load_golden('upenn_tri_calls.npz')
    tp = sfm_b200.TriangulationProcessor(engine=engine)
    for c in range(int(z['num_calls'])):
        g = lambda k: z['c%d_%s' % (c, k)]  # noqa: E731
        init = g('init').copy()
        out = tp.nonlinear_triangulate(init, list(g('projs')), list(g('pairs')), float(g('damping_factor')),
                                       int(g('iteration')))
        assert np.array_equal(init, g('init'))                    # input untouched
        assert close(out, g('out'), TOL), (c, max_rel(out, g('out')))
        assert out.shape == g('out').shape


def test_triangulation_falsy_kwargs_use_instance_defaults(engine, oracle):
    import sfm_b200
    z = load_golden('upenn_tri_calls.npz')
    g = lambda k: z['c4_%s' % k]  # noqa: E731
    tp = sfm_b200.TriangulationProcessor(damping_factor=0.7, iteration=9, engine=engine)
    out = tp.nonlinear_triangulate(g('init'), list(g('projs')), list(g('pairs')), 0, 0)
    ref = oracle.nonlinear_triangulate(g('init'), g('projs'), g('pairs'), 0.7, 9)
    assert close(out, ref, 1e-8)


def test_triangulation_batched_ragged(engine, oracle):
    """1e5 points with 2..6 views each out of 200 projections (config C3 at reduced size)."""
    rng = np.random.default_rng(3)
    P, M = 200, 100000
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
    X0 = X + rng.normal(0, 0.05, X.shape)
    out = engine.triangulate_refine(projs, rowptr, obs_view, uv, X0, 0.5, 20)
    # oracle per track length (it wants the same number of views for every point)
    for k in range(2, 7):
        sel = np.flatnonzero(n == k)[:500]
        for j in sel[:50]:
            o = slice(rowptr[j], rowptr[j + 1])
            init = np.append(X0[j], 1.0).reshape(4, 1)
            pairs = [np.array([[uv[q, 0]], [uv[q, 1]], [1.0]]) for q in range(o.start, o.stop)]
            ref = oracle.nonlinear_triangulate(init, projs[obs_view[o]], pairs, 0.5, 20)
            assert close(out[j], ref[0:3, 0], 1e-8)


def test_pnp_opencv_known_answer(engine):
    """campose_processor.py:1070-1090: lambda 5, 200 iterations; loc_diff 0.012898 vs the OpenCV pose."""
    import sfm_b200
    z = load_golden('opencv_pnp.npz')
    cp = sfm_b200.CamposeProcessor(None, 5, 200, engine=engine)
    rot, loc = cp.nonlinear_estimate_cam_pose_pnp(z['key_2d'], z['tri_3d'], z['k'], z['init_rot'], z['init_loc'], 5, 200)
    assert close(rot, z['out_rot'], TOL), max_rel(rot, z['out_rot'])
    assert close(loc, z['out_loc'], TOL), max_rel(loc, z['out_loc'])
    assert loc.shape == (3, 1) and rot.shape == (3, 3)
    assert abs(np.sqrt(np.sum(np.square(z['loc_truth'] - loc))) - 0.012898) < 1e-5
    rot5, loc5 = cp.nonlinear_estimate_cam_pose_pnp(z['key_2d'], z['tri_3d'], z['k'], z['init_rot'], z['init_loc'], 5, 5)
    assert close(rot5, z['out_rot5'], 1e-9) and close(loc5, z['out_loc5'], 1e-9)


def test_pnp_pipeline_calls(engine):
    import sfm_b200
    z = load_golden('upenn_pnp_calls.npz')
    cp = sfm_b200.CamposeProcessor(None, 5, 300, engine=engine)
    for c in range(int(z['num_calls'])):
        g = lambda k: z['c%d_%s' % (c, k)]  # noqa: E731
        rot, loc = cp.nonlinear_estimate_cam_pose_pnp(g('key_2d'), g('tri_3d'), g('k'), g('init_rot'), g('init_loc'),
                                                      float(g('damping_factor')), int(g('iteration')))
        assert close(rot, g('out_rot'), TOL), (c, max_rel(rot, g('out_rot')))
        assert close(loc, g('out_loc'), TOL), (c, max_rel(loc, g('out_loc')))


def test_pnp_quirk_flags_and_errors(engine, oracle):
    import sfm_b200
    z = load_golden('upenn_pnp_calls.npz')
    g = lambda k: z['c1_%s' % k]  # noqa: E731
    cp = sfm_b200.CamposeProcessor(None, 5, 30, engine=engine)
    for rq in (True, False):
        for dq in (True, False):
            rot, loc = cp.nonlinear_estimate_cam_pose_pnp(g('key_2d'), g('tri_3d'), g('k'), g('init_rot'),
                                                          g('init_loc'), 5, 30, row_packing_quirk=rq,
                                                          dvdc_sign_quirk=dq)
            r_ref, l_ref = oracle.nonlinear_pnp(g('key_2d'), g('tri_3d'), g('k'), g('init_rot'), g('init_loc'), 5, 30,
                                                row_packing_quirk=rq, dvdc_quirk=dq)
            assert close(rot, r_ref, 1e-8) and close(loc, l_ref, 1e-8)
    with pytest.raises(ValueError):
        cp.nonlinear_estimate_cam_pose_pnp(g('key_2d')[:, :5], g('tri_3d'), g('k'), g('init_rot'), g('init_loc'))
    with pytest.raises(ValueError):
        cp.nonlinear_estimate_cam_pose_pnp(g('key_2d'), g('tri_3d'), g('k'), 2.0 * g('init_rot'), g('init_loc'))


def test_pnp_batched_many_views(engine, oracle):
    """Many independent views in one launch (config C3 at reduced size): 300 views x 40..120 points."""
    rng = np.random.default_rng(5)
    P = 300
    n = rng.integers(40, 121, P)
    rowptr = np.zeros(P + 1, dtype=np.int32)
    rowptr[1:] = np.cumsum(n)
    m = int(rowptr[-1])
    rots, locs, xy, X = np.zeros((P, 3, 3)), np.zeros((P, 3)), np.zeros((m, 2)), np.zeros((m, 3))
    from scipy.spatial.transform import Rotation
    for v in range(P):
        R = Rotation.from_rotvec(rng.normal(0, 0.2, 3)).as_matrix()
        C = rng.normal(0, 1, 3)
        pc = np.stack([rng.uniform(-3, 3, n[v]), rng.uniform(-2, 2, n[v]), rng.uniform(5, 15, n[v])], 1)
        o = slice(rowptr[v], rowptr[v + 1])
        X[o] = pc @ R.T + C
        xy[o] = pc[:, 0:2] / pc[:, 2:3] + rng.normal(0, 1e-3, (n[v], 2))
        rots[v] = Rotation.from_rotvec(rng.normal(0, 0.01, 3)).as_matrix() @ R
        locs[v] = C + rng.normal(0, 0.05, 3)
    rot, loc = engine.pnp_refine(rowptr, xy, X, rots, locs, 5.0, 25)
    for v in range(0, P, 17):
        o = slice(rowptr[v], rowptr[v + 1])
        key = np.vstack((xy[o].T, np.ones(n[v])))
        tri = np.vstack((X[o].T, np.ones(n[v])))
        r_ref, l_ref = oracle.nonlinear_pnp(key, tri, np.eye(3), rots[v], locs[v].reshape(3, 1), 5.0, 25)
        assert close(rot[v], r_ref, 1e-8) and close(loc[v], l_ref.reshape(3), 1e-8)


def test_construct_jacobian_matrix_api(engine, oracle):
    import sfm_b200
    case = ba_case_from_npz(load_golden('upenn_ba_call1.npz'))
    cp = sfm_b200.CamposeProcessor(None, 5, 10, engine=engine)
    tp = sfm_b200.TriangulationProcessor(engine=engine)
    t, i, _ = case['vis'][17]
    rot, loc = case['rot'][i], case['loc'][i].reshape(3, 1)
    x = np.append(case['tri_pts'][0:3, t], 1.0).reshape(4, 1)
    jp = cp.construct_jacobian_matrix(rot, loc, x)
    assert jp.shape == (2, 7)
    assert np.max(np.abs(jp - case['jp0'][17]) / np.maximum(1, np.abs(case['jp0'][17]))) < 2e-6
    proj = np.hstack((rot.T, rot.T @ -loc))
    jx = tp.construct_jacobian_matrix(x, [proj], 1)
    assert np.allclose(jx, case['jx0'][17], atol=1e-13)


def test_ba_processor_drop_in(engine):
    """BaProcessor._BaProcessor__execute_bundle_adjustment on the recorded state of the last upenn
    call: views updated through update_cam_pose, tri_pts updated in place, rotations valid."""
    import sfm_b200
    case = ba_case_from_npz(load_golden('upenn_ba_call4.npz'))
    views = [FakeView(case['rot'][i], case['loc'][i], case['k'][i], case['keys'][i]) for i in range(case['V'])]
    tracker = FakeTracker(case['diag'])
    tp = sfm_b200.TriangulationProcessor(engine=engine)
    tp.tri_pts = case['tri_pts'].copy()
    held = tp.tri_pts
    bp = sfm_b200.BaProcessor(FakeViewProcessor(views), tracker, None, tp, None, engine=engine, verbose=False)
    assert bp.iteration == 3 and bp.damping_factor == 5 and bp.filter_size == 10
    bp._BaProcessor__execute_bundle_adjustment()
    assert tp.tri_pts is held                                         # in place (ba_processor.py:416)
    assert close(held, case['out_tri_pts'], TOL)
    for i, v in enumerate(views):
        assert close(v.rot, case['out_rot'][i], TOL) and close(v.loc.reshape(3), case['out_loc'][i], TOL)
        assert sfm_b200.verify_rotation_mat(v.rot)
        sfm_b200.convert_rotation_to_quaternion(v.rot)                # what the next BA call does first
        assert np.allclose(v.cam_proj, v.k @ np.hstack((v.rot.T, v.rot.T @ -v.loc)))
    assert abs(bp.last_result.final_cost - 6.057589389) < 6.1e-6


def test_visibility_on_device_is_bit_exact(engine, oracle):
    for call in (1, 4):
        case = ba_case_from_npz(load_golden('upenn_ba_call%d.npz' % call))
        op, oc, ok, rp = engine.build_visibility(case['diag'], case['N'])
        assert np.array_equal(np.stack([op, oc, ok], 1), case['vis'])
        assert rp[-1] == case['vis'].shape[0] and np.all(np.diff(rp) >= 0)
    z = load_golden('synth_ba_small.npz')
    for c in range(int(z['num_cases'])):
        case = ba_case_from_npz(z, 'c%d_' % c)
        op, oc, ok, rp = engine.build_visibility(case['diag'], case['N'])
        assert np.array_equal(np.stack([op, oc, ok], 1), case['vis'])
    # empty tables
    op, oc, ok, rp = engine.build_visibility([np.full(5, -1), np.full(3, -1)], 4)
    assert op.size == 0 and np.array_equal(rp, np.zeros(5, dtype=np.int32))


# ---- f2: linear triangulation + cheirality on the GPU ------------------------
def test_linear_triangulation_pipeline_calls(engine, oracle):
    """Every linear_triangulate result of the upenn pipeline, recorded from the unmodified reference."""
    import sfm_b200
    z = load_golden('upenn_tri_calls.npz')
    tp = sfm_b200.TriangulationProcessor(engine=engine)
    cp = sfm_b200.CamposeProcessor(None, 5, 10, engine=engine)
    for c in range(5):
        projs, pairs, ref = list(z['c%d_projs' % c]), list(z['c%d_pairs' % c]), z['c%d_init' % c]
        lin = tp.linear_triangulate(projs, pairs)
        assert lin.shape == ref.shape and np.all(lin[3] == 1.0)
        assert np.max(np.abs(lin - ref) / np.maximum(1.0, np.abs(ref))) < 1e-10
        assert cp.evalulate_cam_pose_cheirality(projs[0], projs[1], lin) == oracle.cheirality_indices(projs[0], projs[1], ref)


def test_linear_triangulation_ragged_batch_and_front_flags(engine, oracle):
    """2..6 views per point out of 200 projections, some points behind a camera; vs the oracle's SVD."""
    rng = np.random.default_rng(7)
    P, M = 200, 20000
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
    depth = np.where(rng.random(M) < 0.1, rng.uniform(-14, -6, M), rng.uniform(6, 14, M))   # 10 % behind the cameras
    X = np.stack([first * 0.2 + rng.uniform(-2, 2, M), rng.uniform(-2, 2, M), depth], 1)
    ph = np.einsum('mck,mk->mc', projs[obs_view][:, :, 0:3], X[obs_pt]) + projs[obs_view][:, :, 3]
    uv = ph[:, 0:2] / ph[:, 2:3] + rng.normal(0, 0.5, (rowptr[-1], 2))
    pts, front = engine.triangulate_linear(projs, rowptr, obs_view, uv)
    assert np.array_equal(front, depth > 0)
    for k in range(2, 7):
        sel = np.flatnonzero(n == k)[:200]
        a = rowptr[sel][:, None] + np.arange(k)[None, :]                                      # (S,k) observation ids
        pairs = [np.vstack((uv[a[:, v]].T, np.ones(len(sel)))) for v in range(k)]
        for j, jj in enumerate(sel[:40]):
            ref = oracle.linear_triangulate(list(projs[obs_view[a[j]]]), [p[:, j:j + 1] for p in pairs])
            assert close(pts[jj], ref[0:3, 0], 1e-9), (pts[jj], ref[0:3, 0])


def test_linear_triangulation_matches_opencv(engine):
    """The reference's own check (triangulation_processor.py:456-468): cv2.triangulatePoints."""
    cv2 = pytest.importorskip('cv2')
    z = load_golden('upenn_tri_calls.npz')
    projs, pairs = z['c0_projs'], z['c0_pairs']
    M = pairs.shape[2]
    rowptr = np.arange(M + 1, dtype=np.int32) * 2
    obs_view = np.tile(np.arange(2, dtype=np.int32), M)
    uv = np.stack([pairs[0][0:2].T, pairs[1][0:2].T], axis=1).reshape(-1, 2)
    pts, _ = engine.triangulate_linear(projs, rowptr, obs_view, uv)
    ref = cv2.triangulatePoints(projs[0], projs[1], pairs[0][0:2], pairs[1][0:2])
    ref = (ref[0:3] / ref[3]).T
    assert np.max(np.abs(pts - ref) / np.maximum(1.0, np.abs(ref))) < 1e-8


def test_refiners_reject_bad_indices(engine):
    """rowptr / obs_view are validated on the device: INVALID_ARG instead of an out-of-bounds read."""
    from sfm_b200 import _capi
    rng = np.random.default_rng(11)
    P, M = 5, 40
    projs = np.tile(np.hstack((np.identity(3), np.zeros((3, 1)))), (P, 1, 1))
    rowptr = (2 * np.arange(M + 1)).astype(np.int32)
    obs_view = np.tile(np.array([0, 1], dtype=np.int32), M)
    uv = rng.normal(0, 0.1, (2 * M, 2))
    X0 = np.column_stack([rng.normal(0, 1, M), rng.normal(0, 1, M), rng.uniform(5, 9, M)])

    def variants():
        v = obs_view.copy(); v[17] = P
        yield rowptr, v
        v = obs_view.copy(); v[0] = -2
        yield rowptr, v
        r = rowptr.copy(); r[9], r[10] = r[10], r[9]
        yield r, obs_view
        r = rowptr.copy(); r[-1] += 1
        yield r, obs_view
        r = rowptr.copy(); r[0] = 1
        yield r, obs_view

    for r, v in variants():
        with pytest.raises(_capi.SfmbaError) as err:
            engine.triangulate_refine(projs, r, v, uv, X0, 0.5, 3)
        assert err.value.code == _capi.ERR_INVALID_ARG
        with pytest.raises(_capi.SfmbaError) as err:
            engine.triangulate_linear(projs, r, v, uv)
        assert err.value.code == _capi.ERR_INVALID_ARG
    rot = np.tile(np.identity(3), (M, 1, 1))
    loc = np.zeros((M, 3))
    for r in (rowptr[::-1].copy(), np.where(np.arange(M + 1) == M, 2 * M + 3, rowptr).astype(np.int32)):
        with pytest.raises(_capi.SfmbaError) as err:
            engine.pnp_refine(r, uv, np.column_stack([uv, np.full(2 * M, 6.0)]), rot, loc, 0.5, 2)
        assert err.value.code == _capi.ERR_INVALID_ARG
    # and the valid inputs still run on the same context
    out = engine.triangulate_refine(projs, rowptr, obs_view, uv, X0, 0.5, 3)
    assert out.shape == (M, 3) and np.all(np.isfinite(out))
