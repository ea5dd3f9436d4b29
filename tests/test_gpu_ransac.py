"""GPU parity of the RANSAC initialisers: every hypothesis of the 6-point DLT
PnP (campose_processor.py:485-633) and of the 8-point fundamental matrix (epipolar_processor.py:140-247)
evaluated on the device, against fixtures recorded from the unmodified reference (tests/golden/ransac.npz,
with the sample tables the reference drew) and against the oracle."""
import random

import numpy as np
import pytest

from helpers import load_golden

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('tag', ['clean', 'outliers'])
def test_pnp_ransac_matches_reference(engine, oracle, tag):
    g = load_golden('ransac.npz')
    key, tri, k = g['pnp_%s_key' % tag], g['pnp_tri'], g['pnp_k']
    samples, thr = g['pnp_%s_samples' % tag], float(g['pnp_%s_thr' % tag])
    ora = oracle.pnp_ransac(key, tri, k, samples, thr)
    res = engine.pnp_ransac(key.T, tri.T, k, samples, thr, loc_sign=ora['signs'], want_all=True)
    # every hypothesis: same pose as the reference computes it (numpy SVD) ...
    assert np.max(np.abs(res['models'] - ora['models'])) < 1e-8
    # ... the same integer inlier counts, the same winner and the same inlier list (bit-exact index work)
    assert np.array_equal(res['counts'], ora['counts'])
    assert res['hyp'] == ora['hyp'] and res['count'] == ora['count']
    assert list(np.flatnonzero(res['mask'])) == list(g['pnp_%s_inliers' % tag])
    assert np.max(np.abs(res['rot'] - g['pnp_%s_rot' % tag])) < 1e-9
    assert np.max(np.abs(res['loc'] - g['pnp_%s_loc' % tag])) < 1e-8 * max(1.0, np.abs(g['pnp_%s_loc' % tag]).max())
    assert res['count'] == int(res['mask'].sum())


def test_pnp_ransac_sign_free(engine, oracle):
    """loc_sign = NULL: the geometrically consistent pose (R, C) for every hypothesis, i.e. the reference's pose
    with its LAPACK sign divided out; the winner can only get better."""
    g = load_golden('ransac.npz')
    key, tri, k = g['pnp_outliers_key'], g['pnp_tri'], g['pnp_k']
    samples, thr = g['pnp_outliers_samples'], float(g['pnp_outliers_thr'])
    ora = oracle.pnp_ransac(key, tri, k, samples, thr)
    res = engine.pnp_ransac(key.T, tri.T, k, samples, thr, want_all=True)
    want = ora['models'].copy()
    want[:, 9:12] *= ora['signs'][:, None]
    assert np.max(np.abs(res['models'] - want)) < 1e-8
    assert res['count'] >= ora['count']
    pos = ora['signs'] > 0
    assert np.array_equal(res['counts'][pos], ora['counts'][pos])


def test_campose_processor_linear_pnp_dropin(engine):
    """CamposeProcessor.linear_estimate_cam_pose_pnp: same call, same random stream, same result."""
    import sfm_b200
    g = load_golden('ransac.npz')
    for tag, iters in (('clean', 120), ('outliers', 120)):
        cfg = sfm_b200.RansacConfig(float(g['pnp_%s_thr' % tag]), 0.99, 0.75, 6, iters)     # reseeds `random`
        cp = sfm_b200.CamposeProcessor(cfg, 5, 200, engine=engine)
        inl, rot, loc = cp.linear_estimate_cam_pose_pnp(g['pnp_%s_key' % tag], g['pnp_tri'], g['pnp_k'], cfg)
        after = random.random()
        assert inl == list(g['pnp_%s_inliers' % tag])
        assert np.max(np.abs(rot - g['pnp_%s_rot' % tag])) < 1e-9
        assert np.max(np.abs(loc - g['pnp_%s_loc' % tag])) < 1e-8 * max(1.0, np.abs(g['pnp_%s_loc' % tag]).max())
        assert rot.shape == (3, 3) and loc.shape == (3, 1)
        # the global random stream was consumed exactly as the reference consumes it
        random.seed(-1)
        for _ in range(cfg.iteration):
            random.sample(range(g['pnp_tri'].shape[1]), 6)
        assert random.random() == after
    with pytest.raises(ValueError):
        cp.linear_estimate_cam_pose_pnp(g['pnp_clean_key'][:, :5], g['pnp_tri'][:, :5], g['pnp_k'], cfg)
    with pytest.raises(ValueError):
        cp.linear_estimate_cam_pose_pnp(g['pnp_clean_key'][:, :9], g['pnp_tri'][:, :8], g['pnp_k'], cfg)


@pytest.mark.parametrize('tag', ['clean', 'outliers'])
def test_fund_ransac_matches_reference(engine, oracle, tag):
    g = load_golden('ransac.npz')
    pairs, tl, tr = oracle.fund_normalize(g['fund_%s_left' % tag], g['fund_%s_right' % tag])
    samples, thr = g['fund_%s_samples' % tag], float(g['fund_%s_thr' % tag])
    ora = oracle.fund_ransac(pairs, samples, thr)
    res = engine.fund_ransac(pairs, samples, thr, want_all=True)
    scale = np.max(np.abs(ora['models']), axis=1, keepdims=True)
    assert np.max(np.abs(res['models'] - ora['models']) / scale) < 1e-8
    assert np.array_equal(res['counts'], ora['counts'])
    assert res['hyp'] == ora['hyp'] and res['count'] == ora['count']
    assert list(np.flatnonzero(res['mask'])) == list(g['fund_%s_inliers' % tag])
    f = oracle.fund_denormalize(res['F'], tl, tr)
    assert np.max(np.abs(f - g['fund_%s_F' % tag])) < 1e-8 * np.max(np.abs(g['fund_%s_F' % tag]))


def test_epipolar_processor_dropin(engine):
    import sfm_b200
    g = load_golden('ransac.npz')
    for tag, iters in (('clean', 60), ('outliers', 60)):
        n = g['fund_%s_left' % tag].shape[0]
        left, right = sfm_b200.KeyPt(n), sfm_b200.KeyPt(n)
        left[0:2, :] = g['fund_%s_left' % tag].T
        right[0:2, :] = g['fund_%s_right' % tag].T
        cfg = sfm_b200.RansacConfig(float(g['fund_%s_thr' % tag]), 0.99, 0.75, 8, iters)
        ep = sfm_b200.EpipolarProcessor(cfg, engine=engine)
        inl = ep.determine_fundamental_mat([left, right], cfg)
        assert inl == list(g['fund_%s_inliers' % tag])
        assert np.max(np.abs(ep.fund_mat - g['fund_%s_F' % tag])) < 1e-8 * np.max(np.abs(g['fund_%s_F' % tag]))
    # exactly eight pairs: no RANSAC, epipolar_processor.py:216-218 and the known answer of :274-299
    d8 = g['fund_eight_data']
    left, right = sfm_b200.KeyPt(8), sfm_b200.KeyPt(8)
    left[0:2, :] = d8[:, 0:2].T
    right[0:2, :] = d8[:, 2:4].T
    cfg = sfm_b200.RansacConfig(1e-3, 0.99, 0.75, 8, 200)
    ep = sfm_b200.EpipolarProcessor(cfg, engine=engine)
    assert ep.determine_fundamental_mat([left, right], cfg) == list(range(8))
    assert np.max(np.abs(ep.fund_mat - g['fund_eight_F'])) < 1e-7 * np.max(np.abs(g['fund_eight_F']))
    with pytest.raises(ValueError):
        ep.determine_fundamental_mat([left[:, :7], right[:, :7]], cfg)


def test_ransac_large_synthetic(engine, oracle):
    """50 000 points x 1 000 hypotheses of an exact synthetic pose: sign-free hypotheses drawn from inliers
    recover the pose and count every inlier; a sample of hypotheses is checked against the oracle."""
    rng = np.random.default_rng(11)
    n, H = 50000, 1000
    k = np.array([[800.0, 0, 320], [0, 800, 240], [0, 0, 1]])
    ang = 0.3
    rot = np.array([[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]])
    loc = np.array([[0.5], [-0.2], [0.1]])
    X = np.vstack([rng.uniform(-2, 2, (2, n)), rng.uniform(4, 9, (1, n)), np.ones((1, n))])
    p = k @ np.hstack((rot.T, rot.T @ -loc)) @ X
    key = p / p[2]
    bad = rng.random(n) < 0.3
    key[0:2, bad] += rng.normal(0, 40, (2, int(bad.sum())))
    samples = np.array([rng.choice(n, 6, replace=False) for _ in range(H)], dtype=np.int32)
    res = engine.pnp_ransac(key.T, X.T, k, samples, 1.0, want_all=True)
    assert res['count'] == int(res['mask'].sum()) == int(res['counts'].max())
    assert res['hyp'] == int(np.argmax(res['counts']))            # first maximum
    assert res['count'] >= int((~bad).sum())
    assert np.max(np.abs(res['rot'] - rot)) < 1e-6 and np.max(np.abs(res['loc'] - loc)) < 1e-5
    sub = np.arange(0, H, 97)
    ora = oracle.pnp_ransac(key, X, k, samples[sub], 1.0)
    want = ora['models'].copy()
    want[:, 9:12] *= ora['signs'][:, None]
    ok = np.all(np.isfinite(want), axis=1)
    assert np.max(np.abs(res['models'][sub][ok] - want[ok])) < 1e-6


def test_ransac_edge_cases(engine):
    """No hypotheses, no inliers, bad indices, degenerate samples."""
    import sfm_b200
    g = load_golden('ransac.npz')
    key, tri, k = g['pnp_clean_key'], g['pnp_tri'], g['pnp_k']
    # no hypotheses: the reference's initial best (identity, zero, no inliers), campose_processor.py:524-527
    res = engine.pnp_ransac(key.T, tri.T, k, np.zeros((0, 6), dtype=np.int32), 8.0)
    assert res['hyp'] == -1 and res['count'] == 0 and not res['mask'].any()
    assert np.array_equal(res['rot'], np.identity(3)) and np.array_equal(res['loc'], np.zeros((3, 1)))
    # a threshold nothing passes: same outcome with hypotheses present
    res = engine.pnp_ransac(key.T, tri.T, k, g['pnp_clean_samples'][:10], 1e-12)
    assert res['hyp'] == -1 and res['count'] == 0 and not res['mask'].any()
    with pytest.raises(ValueError):
        engine.pnp_ransac(key.T, tri.T, k, np.full((1, 6), key.shape[1], dtype=np.int32), 8.0)
    with pytest.raises(ValueError):
        engine.fund_ransac(np.zeros((10, 4)), np.full((1, 8), 10, dtype=np.int32), 1.0)
    # eight identical pairs: f__ cannot have rank 2 -> the reference raises (epipolar_processor.py:179-183)
    pairs = np.tile(np.array([[0.1, 0.2, 0.3, 0.4]]), (12, 1))
    res = engine.fund_ransac(pairs, np.arange(8, dtype=np.int32).reshape(1, 8), 1.0)
    assert res['flags'] & 16 or not np.all(np.isfinite(res['F']))
    cfg = sfm_b200.RansacConfig(1e-3, 0.99, 0.75, 8, 1, is_use_seed=False)
    ep = sfm_b200.EpipolarProcessor(cfg, engine=engine)
    with pytest.raises(ValueError):
        ep._check_rank(engine.fund_ransac(np.tile(np.array([[0.1, 0.2, 0.3, 0.4]]), (9, 1)),
                                          np.arange(8, dtype=np.int32).reshape(1, 8), 1.0, want_all=True))


def test_ransac_c_abi_rejects_bad_sample_index(engine, oracle):
    """The C-ABI itself (not only the Python wrapper) refuses a sample index outside [0, n)."""
    import ctypes as C
    from sfm_b200 import _capi
    from sfm_b200.engine import _ptr
    g = load_golden('ransac.npz')
    key = np.ascontiguousarray(g['pnp_clean_key'].T)
    tri = np.ascontiguousarray(g['pnp_tri'].T)
    k = np.ascontiguousarray(g['pnp_k'], dtype=np.float64)
    n = key.shape[0]
    for bad in (n, -1):
        samples = np.ascontiguousarray(g['pnp_clean_samples'][:4], dtype=np.int32).copy()
        samples[2, 3] = bad
        rot, loc, best = np.zeros(9), np.zeros(3), np.zeros(2, dtype=np.int32)
        mask, flags = np.zeros(n, dtype=np.uint8), C.c_uint32(0)
        rc = engine.lib.sfmba_pnp_ransac_host(engine.ctx, n, 4, _ptr(key), _ptr(tri), _ptr(k), _ptr(samples), None, 8.0,
                                              _ptr(rot), _ptr(loc), _ptr(best), _ptr(mask), None, None, C.byref(flags))
        assert rc == _capi.ERR_INVALID_ARG
        pairs = np.ascontiguousarray(oracle.fund_normalize(g['fund_clean_left'], g['fund_clean_right'])[0], dtype=np.float64)
        fs = np.ascontiguousarray(g['fund_clean_samples'][:4], dtype=np.int32).copy()
        fs[1, 7] = bad if bad < 0 else pairs.shape[0]
        fm = np.zeros(9)
        fmask = np.zeros(pairs.shape[0], dtype=np.uint8)
        rc = engine.lib.sfmba_fund_ransac_host(engine.ctx, pairs.shape[0], 4, _ptr(pairs), _ptr(fs), 1e-3,
                                               _ptr(fm), _ptr(best), _ptr(fmask), None, None, C.byref(flags))
        assert rc == _capi.ERR_INVALID_ARG
