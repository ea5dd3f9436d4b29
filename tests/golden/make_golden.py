#!/usr/bin/env python3
"""Generate the golden fixtures under tests/golden/ by running the UNMODIFIED reference.

TEST INFRASTRUCTURE.  Runs only in the build container (needs /root/reference, which does
not exist on the GPU box).  Nothing from the reference is copied: it is imported from where
it lies and its hot-path methods are wrapped so that their inputs and outputs are recorded.

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py [--skip-pipeline] [--skip-pnp]

Fixtures written (all float64 unless noted):
  upenn_ba_call{1..4}.npz   inputs/outputs of the four BaProcessor.__execute_bundle_adjustment
                            calls of `python3 ba_processor.py` (ba_processor.py:274-440), the
                            is_visible() loop-order index list (ba_processor.py:304-310) and the
                            literal per-observation Jacobians of the first iteration
                            (campose_processor.py:462-482, triangulation_processor.py:237-271)
  upenn_final.npz           end state of the pipeline + the reference's stored results/*.pkl
  upenn_tri_calls.npz       every nonlinear_triangulate call of the pipeline (a18)
  upenn_pnp_calls.npz       every nonlinear_estimate_cam_pose_pnp call of the pipeline (a19)
  opencv_pnp.npz            campose_processor.py:1070-1090 nonlinear PnP known answer
  tri_inline.npz            triangulation_processor.py:415-473 inline known answer
  synth_ba_small.npz        random small BA problems pushed through the literal reference
  ransac.npz                f3: linear_estimate_cam_pose_pnp and determine_fundamental_mat on the OpenCV test
                            data with the sample tables the reference drew (campose_processor.py:485-633,
                            epipolar_processor.py:23-265)
"""
import argparse
import os
import sys
import types
import warnings

import numpy as np

REF = '/root/reference'
HERE = os.path.dirname(os.path.abspath(__file__))

warnings.filterwarnings('ignore')
sys.dont_write_bytecode = True


def _install_matplotlib_stub():
    """ba_processor.py:11-12 imports matplotlib at module scope; it is not installed here."""
    class _Anything(types.ModuleType):
        def __getattr__(self, name):
            if name.startswith('__'):
                raise AttributeError(name)
            return _Callable()

    class _Callable:
        def __call__(self, *a, **k):
            return _Callable()

        def __getattr__(self, name):
            if name.startswith('__'):
                raise AttributeError(name)
            return _Callable()

    for name in ('matplotlib', 'matplotlib.pyplot', 'matplotlib.markers'):
        sys.modules[name] = _Anything(name)
    sys.modules['matplotlib'].pyplot = sys.modules['matplotlib.pyplot']
    sys.modules['matplotlib'].markers = sys.modules['matplotlib.markers']


def _import_reference():
    _install_matplotlib_stub()
    sys.path.insert(0, REF)
    import ba_processor
    import campose_processor
    import key_tracker
    import triangulation_processor
    import utils
    import view_processor
    import epipolar_processor
    return types.SimpleNamespace(ba=ba_processor, cp=campose_processor, kt=key_tracker,
                                 tp=triangulation_processor, utils=utils, vp=view_processor,
                                 ep=epipolar_processor)


def _snapshot_ba_inputs(bp):
    views = bp.view_processor.view_list
    snap = {
        'rot': np.stack([v.rot.copy() for v in views]),
        'loc': np.stack([v.loc.copy().reshape(3) for v in views]),
        'k': np.stack([v.k.copy() for v in views]),
        'tri_pts': np.array(bp.tri_processor.tri_pts, dtype=np.float64).copy(),
    }
    for i, v in enumerate(views):
        snap['keys_%d' % i] = np.array([kp.pt for kp in v.key_pts], dtype=np.float32)
        snap['diag_%d' % i] = bp.key_tracker.track_list[i].table[i, :].copy()
    return snap


def _visibility_loop_order(bp):
    """The literal (tri_idx, view_idx, key_idx) list of ba_processor.py:304-310."""
    out = []
    n = bp.tri_processor.tri_pts.shape[1]
    v = len(bp.view_processor.view_list)
    for t in range(n):
        for i in range(v):
            k = bp.key_tracker.is_visible(i, t)
            if k != -1:
                out.append((t, i, k))
    return np.array(out, dtype=np.int64).reshape(-1, 3)


def _literal_jacobians(ref, bp, snap, vis):
    """Per-observation 2x7 / 2x3 blocks and b, f at the *initial* parameters, computed with the
    reference's own functions exactly as ba_processor.py:317-349 calls them."""
    m = vis.shape[0]
    jp = np.zeros((m, 2, 7))
    jx = np.zeros((m, 2, 3))
    b = np.zeros((m, 2))
    f = np.zeros((m, 2))
    for o, (t, i, k) in enumerate(vis):
        rot = snap['rot'][i]
        qua = ref.utils.convert_rotation_to_quaternion(rot)
        rot = ref.utils.convert_quaternion_to_rotation(qua)
        loc = snap['loc'][i].reshape(3, 1)
        tri_pt = np.append(snap['tri_pts'][0:3, t:t + 1], [[1.0]], axis=0)
        proj = np.hstack((rot.T, rot.T @ -loc))
        jp[o] = bp.campose_processor.construct_jacobian_matrix(rot, loc, tri_pt)
        jx[o] = bp.tri_processor.construct_jacobian_matrix(tri_pt, [proj], 1)
        key = snap['keys_%d' % i][k]
        key_pt = np.array([[key[0], key[1], 1.0]]).T
        key_pt = np.linalg.inv(snap['k'][i]) @ key_pt
        key_pt /= key_pt[2]
        b[o] = key_pt[0:2, 0]
        pp = proj @ tri_pt
        pp /= pp[2]
        f[o] = pp[0:2, 0]
    return jp, jx, b, f


def run_pipeline(ref):
    import cv2 as cv
    K = np.array([[568.996140852, 0, 643.21055941],
                  [0, 568.988362396, 477.982801038],
                  [0, 0, 1]])
    RC = ref.utils.RansacConfig
    ransac_kt = RC(inlier_threshold=1e-2, subset_confidence=0.99, sample_confidence=0.75, sample_num=8, iteration=200)
    ransac_ep = RC(inlier_threshold=1e-3, subset_confidence=0.99, sample_confidence=0.75, sample_num=8, iteration=300)
    ransac_cp = RC(inlier_threshold=8.0, subset_confidence=0.99, sample_confidence=0.75, sample_num=8, iteration=300)
    vp = ref.vp.ViewProcessor('sift')
    kt = ref.kt.KeyTracker('sift', False, True, False, ransac_kt)
    ep = ref.ep.EpipolarProcessor(ransac_ep)
    tp = ref.tp.TriangulationProcessor()
    cp = ref.cp.CamposeProcessor(ransac_cp, 5, 300)
    bp = ref.ba.BaProcessor(vp, kt, ep, tp, cp)

    ba_calls = []
    orig_ba = bp._BaProcessor__execute_bundle_adjustment

    def wrapped_ba():
        snap = _snapshot_ba_inputs(bp)
        vis = _visibility_loop_order(bp)
        jp, jx, b, f = _literal_jacobians(ref, bp, snap, vis)
        orig_ba()
        views = bp.view_processor.view_list
        snap.update({
            'vis': vis, 'jp0': jp, 'jx0': jx, 'b0': b, 'f0': f,
            'out_rot': np.stack([v.rot.copy() for v in views]),
            'out_loc': np.stack([v.loc.copy().reshape(3) for v in views]),
            'out_tri_pts': np.array(bp.tri_processor.tri_pts, dtype=np.float64).copy(),
            'iteration': np.int64(bp.iteration), 'damping_factor': np.float64(bp.damping_factor),
        })
        ba_calls.append(snap)
        print('  captured BA call %d: V=%d N=%d m=%d' % (len(ba_calls), len(views),
              snap['tri_pts'].shape[1], vis.shape[0]), flush=True)

    bp._BaProcessor__execute_bundle_adjustment = wrapped_ba

    tri_calls = []
    orig_tri = tp.nonlinear_triangulate

    def wrapped_tri(init_3d_pts, projs, matched_pairs, damping_factor=None, iteration=None):
        out = orig_tri(init_3d_pts, projs, matched_pairs, damping_factor, iteration)
        tri_calls.append({
            'init': np.array(init_3d_pts, dtype=np.float64).copy(),
            'projs': np.stack([np.array(p, dtype=np.float64) for p in projs]),
            'pairs': np.stack([np.array(p, dtype=np.float64) for p in matched_pairs]),
            'damping_factor': np.float64(damping_factor if damping_factor else tp.damping_factor),
            'iteration': np.int64(iteration if iteration else tp.iteration),
            'out': np.array(out, dtype=np.float64).copy(),
        })
        return out

    tp.nonlinear_triangulate = wrapped_tri

    pnp_calls = []
    orig_pnp = cp.nonlinear_estimate_cam_pose_pnp

    def wrapped_pnp(key_2d_pts, tri_3d_pts, intrinsic_mat, init_rot, init_loc, damping_factor=None, iteration=None):
        rec = {
            'key_2d': np.array(key_2d_pts, dtype=np.float64).copy(),
            'tri_3d': np.array(tri_3d_pts, dtype=np.float64).copy(),
            'k': np.array(intrinsic_mat, dtype=np.float64).copy(),
            'init_rot': np.array(init_rot, dtype=np.float64).copy(),
            'init_loc': np.array(init_loc, dtype=np.float64).copy(),
            'damping_factor': np.float64(damping_factor if damping_factor else cp.damping_factor),
            'iteration': np.int64(iteration if iteration else cp.iteration),
        }
        rot, loc = orig_pnp(key_2d_pts, tri_3d_pts, intrinsic_mat, init_rot, init_loc, damping_factor, iteration)
        rec['out_rot'] = np.array(rot).copy()
        rec['out_loc'] = np.array(loc).copy()
        pnp_calls.append(rec)
        return rot, loc

    cp.nonlinear_estimate_cam_pose_pnp = wrapped_pnp

    files = ['image%07d.bmp' % i for i in range(1, 7)]
    for name in files:
        print('processing', name, flush=True)
        img = cv.imread(os.path.join(REF, 'test_dataset', 'upenn', name))
        bp.process(img, K)

    for c, snap in enumerate(ba_calls):
        np.savez_compressed(os.path.join(HERE, 'upenn_ba_call%d.npz' % (c + 1)), **snap)

    import pickle
    vpk = pickle.load(open(os.path.join(REF, 'test_dataset/upenn/results/view_pose.pkl'), 'rb'))
    tpk = pickle.load(open(os.path.join(REF, 'test_dataset/upenn/results/tri_pts.pkl'), 'rb'))
    views = bp.view_processor.view_list
    np.savez_compressed(os.path.join(HERE, 'upenn_final.npz'),
                        loc=np.stack([v.loc.reshape(3) for v in views]),
                        rot=np.stack([v.rot for v in views]),
                        tri_pts=np.array(bp.tri_processor.tri_pts, dtype=np.float64),
                        stored_loc=np.asarray(vpk[0]), stored_rot=np.asarray(vpk[1]),
                        stored_tri_pts=np.asarray(tpk[0]))

    flat = {}
    for c, rec in enumerate(tri_calls):
        for key, val in rec.items():
            flat['c%d_%s' % (c, key)] = val
    flat['num_calls'] = np.int64(len(tri_calls))
    np.savez_compressed(os.path.join(HERE, 'upenn_tri_calls.npz'), **flat)

    flat = {}
    for c, rec in enumerate(pnp_calls):
        for key, val in rec.items():
            flat['c%d_%s' % (c, key)] = val
    flat['num_calls'] = np.int64(len(pnp_calls))
    np.savez_compressed(os.path.join(HERE, 'upenn_pnp_calls.npz'), **flat)
    print('pipeline fixtures written')


def run_opencv_pnp(ref):
    """campose_processor.py:979-1090 without the OpenCV comparison."""
    d = os.path.join(REF, 'test_dataset', 'opencv')
    ref_k = np.load(d + '/ess_intrinsic_mat.npy')
    pts3 = np.load(d + '/pnp_points_3d.npy').T
    ones = np.ones((1, pts3.shape[1]))
    pts3h = np.vstack((pts3, ones))
    pts2h = np.vstack((np.load(d + '/pnp_points_2d.npy').T, ones))
    rot_truth = np.load(d + '/pnp_rotation.npy').T
    tra_truth = np.load(d + '/pnp_translation.npy')
    loc_truth = rot_truth @ -tra_truth
    cfg = ref.utils.RansacConfig(inlier_threshold=8.0, subset_confidence=0.99, sample_confidence=0.75,
                                 sample_num=6, iteration=300)
    cp = ref.cp.CamposeProcessor(cfg, 5, 200)
    inl, rot, loc = cp.linear_estimate_cam_pose_pnp(pts2h, pts3h, ref_k, cfg)
    key2 = pts2h[:, inl]
    tri3 = pts3h[:, inl]
    out_rot, out_loc = cp.nonlinear_estimate_cam_pose_pnp(key2, tri3, ref_k, rot, loc, 5, 200)
    # a short run as well so the parity test can also pin early iterations
    out_rot5, out_loc5 = cp.nonlinear_estimate_cam_pose_pnp(key2, tri3, ref_k, rot, loc, 5, 5)
    np.savez_compressed(os.path.join(HERE, 'opencv_pnp.npz'), key_2d=key2, tri_3d=tri3, k=ref_k,
                        init_rot=rot, init_loc=loc, out_rot=out_rot, out_loc=out_loc,
                        out_rot5=out_rot5, out_loc5=out_loc5,
                        rot_truth=rot_truth, loc_truth=loc_truth,
                        damping_factor=np.float64(5), iteration=np.int64(200))
    print('opencv_pnp: loc_diff %.6f' % np.sqrt(np.sum(np.square(loc_truth - out_loc))))


def run_ransac(ref):
    """f3: the two RANSAC initialisers on the reference's own OpenCV test data, with the sample table the
    reference drew (random.sample replayed from the same generator state)."""
    import random
    d = os.path.join(REF, 'test_dataset', 'opencv')
    # ---- 6-point DLT PnP (campose_processor.py:249-305, 485-633)
    ref_k = np.load(d + '/ess_intrinsic_mat.npy')
    pts3 = np.load(d + '/pnp_points_3d.npy').T
    ones = np.ones((1, pts3.shape[1]))
    pts3h = np.vstack((pts3, ones))
    pts2h = np.vstack((np.load(d + '/pnp_points_2d.npy').T, ones))
    # a perturbed copy: every 5th point is an outlier, so the inlier sets differ between hypotheses
    rng = np.random.default_rng(7)
    pts2o = pts2h.copy()
    out = np.arange(0, pts2o.shape[1], 5)
    pts2o[0:2, out] += rng.normal(0, 60.0, size=(2, out.size))
    res = {}
    for tag, key, thr, iters in (('clean', pts2h, 8.0, 120), ('outliers', pts2o, 30.0, 120)):
        cfg = ref.utils.RansacConfig(inlier_threshold=thr, subset_confidence=0.99, sample_confidence=0.75,
                                     sample_num=6, iteration=iters)
        state = random.getstate()
        cp = ref.cp.CamposeProcessor(cfg, 5, 200)
        inl, rot, loc = cp.linear_estimate_cam_pose_pnp(key, pts3h, ref_k, cfg)
        random.setstate(state)
        samples = np.array([random.sample(range(key.shape[1]), 6) for _ in range(cfg.iteration)], dtype=np.int32)
        res['pnp_%s_key' % tag] = key
        res['pnp_%s_samples' % tag] = samples
        res['pnp_%s_inliers' % tag] = np.array(inl, dtype=np.int32)
        res['pnp_%s_rot' % tag] = rot
        res['pnp_%s_loc' % tag] = loc
        res['pnp_%s_thr' % tag] = np.float64(thr)
        print('pnp_ransac %s: %d hypotheses, %d inliers' % (tag, cfg.iteration, len(inl)))
    res['pnp_tri'] = pts3h
    res['pnp_k'] = ref_k
    # ---- 8-point fundamental matrix (epipolar_processor.py:23-56, 96-265)
    p1 = np.load(d + '/ess_pixel_pt1.npy')[:, 0:2]
    p2 = np.load(d + '/ess_pixel_pt2.npy')[:, 0:2]
    p2o = p2.copy()
    out = np.arange(0, p2o.shape[0], 4)
    p2o[out] += rng.normal(0, 25.0, size=(out.size, 2))
    ep_mod = ref.ep
    for tag, right, thr, iters in (('clean', p2, 1e-3, 60), ('outliers', p2o, 1e-1, 60)):
        left_pts = ref.utils.KeyPt(p1.shape[0])
        right_pts = ref.utils.KeyPt(p1.shape[0])
        left_pts[0:2, :] = p1.T
        right_pts[0:2, :] = right.T
        cfg = ref.utils.RansacConfig(inlier_threshold=thr, subset_confidence=0.99, sample_confidence=0.75,
                                     sample_num=8, iteration=iters)
        state = random.getstate()
        ep = ep_mod.EpipolarProcessor(cfg)
        inl = ep.determine_fundamental_mat([left_pts, right_pts], cfg)
        random.setstate(state)
        samples = np.array([random.sample(range(p1.shape[0]), 8) for _ in range(cfg.iteration)], dtype=np.int32)
        res['fund_%s_left' % tag] = p1
        res['fund_%s_right' % tag] = right
        res['fund_%s_samples' % tag] = samples
        res['fund_%s_inliers' % tag] = np.array(inl, dtype=np.int32)
        res['fund_%s_F' % tag] = ep.fund_mat
        res['fund_%s_thr' % tag] = np.float64(thr)
        print('fund_ransac %s: %d hypotheses, %d inliers' % (tag, cfg.iteration, len(inl)))
    # the 8-pair known answer of epipolar_processor.py:274-299 (rows == 8: no RANSAC)
    data8 = np.array([[580, 2362, 492, 1803], [2050, 2097, 1381, 1956], [2558, 2174, 1544, 2115],
                      [1395, 1970, 1166, 1752], [2490, 3003, 466, 2440], [3368, 1622, 3320, 2011],
                      [2183, 1500, 2471, 1621], [1972, 1775, 1674, 1736]], dtype=float)
    left_pts = ref.utils.KeyPt(8)
    right_pts = ref.utils.KeyPt(8)
    left_pts[0:2, :] = data8[:, 0:2].T
    right_pts[0:2, :] = data8[:, 2:4].T
    cfg = ref.utils.RansacConfig(inlier_threshold=1e-3, subset_confidence=0.99, sample_confidence=0.75,
                                 sample_num=8, iteration=200)
    ep = ep_mod.EpipolarProcessor(cfg)
    inl = ep.determine_fundamental_mat([left_pts, right_pts], cfg)
    res['fund_eight_data'] = data8
    res['fund_eight_F'] = ep.fund_mat
    res['fund_eight_inliers'] = np.array(inl, dtype=np.int32)
    np.savez_compressed(os.path.join(HERE, 'ransac.npz'), **res)


def run_tri_inline(ref):
    p1 = np.array([[5.010e+03, 0.000e+00, 3.600e+02, 0.000e+00],
                   [0.000e+00, 5.010e+03, 6.400e+02, 0.000e+00],
                   [0.000e+00, 0.000e+00, 1.000e+00, 0.000e+00]])
    p2 = np.array([[5.037e+03, -9.611e+01, -1.756e+03, 4.284e+03],
                   [2.148e+02, 5.354e+03, 1.918e+02, 8.945e+02],
                   [3.925e-01, 7.092e-02, 9.169e-01, 4.930e-01]])
    p3 = np.array([[5.217e+03, 2.246e+02, 2.366e+03, -3.799e+03],
                   [-5.734e+02, 5.669e+03, 8.233e+02, -2.567e+02],
                   [-3.522e-01, -5.839e-02, 9.340e-01, 6.459e-01]])
    x = [np.array([[274.128, 624.409, 1.0]]).T, np.array([[239.571, 533.568, 1.0]]).T,
         np.array([[297.574, 549.260, 1.0]]).T]
    tp = ref.tp.TriangulationProcessor()
    pairs = []
    for xi in x[:2]:
        kp = ref.utils.KeyPt(1)
        kp[:, :] = xi
        pairs.append(kp)
    lin = tp.linear_triangulate([p1, p2], pairs)
    whole = tp.nonlinear_triangulate(lin, [p1, p2], pairs, 0.5, 300)
    # three-view variant: the reference's own three-view self-test is commented out
    # (triangulation_processor.py:523-583) but nonlinear_triangulate itself supports n views.
    pairs3 = []
    for xi in x:
        kp = ref.utils.KeyPt(1)
        kp[:, :] = xi
        pairs3.append(kp)
    whole3 = tp.nonlinear_triangulate(lin, [p1, p2, p3], pairs3, 0.5, 300)
    jac = tp.construct_jacobian_matrix(np.array(lin), [p1, p2, p3], 3)
    np.savez_compressed(os.path.join(HERE, 'tri_inline.npz'), projs=np.stack([p1, p2, p3]),
                        pairs=np.stack([np.array(p) for p in pairs3]), linear=np.array(lin),
                        out2=np.array(whole), out3=np.array(whole3), jac3=jac)
    print('tri_inline:', np.array(lin).T, np.array(whole).T)


class _FakeKey:
    def __init__(self, pt):
        self.pt = (float(pt[0]), float(pt[1]))


def run_synth_small(ref):
    """Small random BA problems through the literal BaProcessor.__execute_bundle_adjustment.
    Views / tracks are built by hand (no images): the literal method only touches
    view.rot/.loc/.k/.key_pts[k].pt/.update_cam_pose, tri_pts and key_tracker.is_visible."""
    rng = np.random.default_rng(7)
    out = {}
    cases = [dict(V=4, N=40, seed=1, it=3, lam=5.0), dict(V=7, N=90, seed=2, it=4, lam=1.0),
             dict(V=3, N=25, seed=3, it=2, lam=5.0)]
    for ci, case in enumerate(cases):
        V, N = case['V'], case['N']
        rng = np.random.default_rng(case['seed'])
        K = np.array([[569.0, 0, 640.0], [0, 569.0, 480.0], [0, 0, 1]])
        views = []
        rots, locs = [], []
        for i in range(V):
            ang = rng.normal(0, 0.08, 3)
            from scipy.spatial.transform import Rotation
            R = Rotation.from_rotvec(ang).as_matrix()
            C = np.array([[i * 1.0 + rng.normal(0, 0.1)], [rng.normal(0, 0.1)], [rng.normal(0, 0.1)]])
            rots.append(R)
            locs.append(C)
        X = np.vstack((rng.uniform(-2, V + 1, N), rng.uniform(-3, 3, N), rng.uniform(5, 15, N), np.ones(N)))
        # visibility: every point seen by 2..V views; key index 0 deliberately used for a few
        # points (is_visible index-0 quirk, key_tracker.py:198-204) and one point unobserved
        tables = []
        keys = []
        for i in range(V):
            keys.append([])
        diag = [[] for _ in range(V)]
        for t in range(N):
            if t == N - 1:
                continue  # unobserved point: D = lambda I (ba_processor.py:359)
            nv = int(rng.integers(2, V + 1))
            vs = np.sort(rng.choice(V, nv, replace=False))
            for i in vs:
                p = rots[i].T @ (X[0:3, t:t + 1] - locs[i])
                uv = (K @ (p / p[2]))[0:2, 0] + rng.normal(0, 0.5, 2)
                keys[i].append(np.float32(uv))
                diag[i].append(t)
        view_objs = []
        track_list = []
        for i in range(V):
            # shuffle the key order, add some unused keys
            nk = len(keys[i])
            extra = int(rng.integers(1, 6))
            perm = rng.permutation(nk + extra)
            kp = [None] * (nk + extra)
            dg = np.full(nk + extra, -1, dtype=int)
            for src in range(nk + extra):
                dst = perm[src]
                if src < nk:
                    kp[dst] = _FakeKey(keys[i][src])
                    dg[dst] = diag[i][src]
                else:
                    kp[dst] = _FakeKey(rng.uniform(0, 1000, 2).astype(np.float32))
            # duplicate assignment: one extra key mapped to an existing point (several matches,
            # smallest index wins; key_tracker.py:199-201)
            if nk > 3 and extra > 0:
                free = np.where(dg == -1)[0]
                dg[free[0]] = dg[np.where(dg != -1)[0][0]]
            view = ref.vp.View(None, i, K.copy(), kp, None)
            view.update_cam_pose(rots[i] + 0.0, locs[i] + 0.0)
            view_objs.append(view)
            tr = ref.kt.KeyTrack(V, nk + extra, i)
            tr.table[i, :] = dg
            track_list.append(tr)
        # perturbed initial guess
        Xn = X.copy()
        Xn[0:3] += rng.normal(0, 0.05, (3, N)) * X[2]
        for i, view in enumerate(view_objs):
            from scipy.spatial.transform import Rotation
            dR = Rotation.from_rotvec(rng.normal(0, 0.01, 3)).as_matrix()
            view.update_cam_pose(dR @ rots[i], locs[i] + rng.normal(0, 0.05, (3, 1)))
        vp = types.SimpleNamespace(view_list=view_objs)
        kt = ref.kt.KeyTracker('sift', False, True, False, None)
        kt.track_list = track_list
        tp = ref.tp.TriangulationProcessor()
        tp.tri_pts = Xn
        cp = ref.cp.CamposeProcessor(None, 5, 10)
        bp = ref.ba.BaProcessor(vp, kt, None, tp, cp, iteration=case['it'], damping_factor=case['lam'])
        snap = _snapshot_ba_inputs(bp)
        vis = _visibility_loop_order(bp)
        jp, jx, b, f = _literal_jacobians(ref, bp, snap, vis)
        import io
        import contextlib
        with contextlib.redirect_stdout(io.StringIO()):
            bp._BaProcessor__execute_bundle_adjustment()
        snap.update({'vis': vis, 'jp0': jp, 'jx0': jx, 'b0': b, 'f0': f,
                     'out_rot': np.stack([v.rot.copy() for v in view_objs]),
                     'out_loc': np.stack([v.loc.copy().reshape(3) for v in view_objs]),
                     'out_tri_pts': np.array(tp.tri_pts, dtype=np.float64).copy(),
                     'iteration': np.int64(case['it']), 'damping_factor': np.float64(case['lam'])})
        for key, val in snap.items():
            out['c%d_%s' % (ci, key)] = val
        out['c%d_V' % ci] = np.int64(V)
        print('  synth case %d: V=%d N=%d m=%d' % (ci, V, N, vis.shape[0]), flush=True)
    out['num_cases'] = np.int64(len(cases))
    np.savez_compressed(os.path.join(HERE, 'synth_ba_small.npz'), **out)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--skip-pipeline', action='store_true')
    ap.add_argument('--skip-pnp', action='store_true')
    ap.add_argument('--skip-synth', action='store_true')
    ap.add_argument('--only-ransac', action='store_true')
    args = ap.parse_args()
    ref = _import_reference()
    run_ransac(ref)
    if args.only_ransac:
        return
    run_tri_inline(ref)
    if not args.skip_synth:
        run_synth_small(ref)
    if not args.skip_pnp:
        run_opencv_pnp(ref)
    if not args.skip_pipeline:
        run_pipeline(ref)


if __name__ == '__main__':
    main()
