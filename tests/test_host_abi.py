"""CPU-side checks: the C-ABI library loads and exports every symbol include/sfmba.h declares, the
product path fails loudly without a GPU (no CPU fallback), and the host-side packing / index /
utility code agrees with the oracle.  No compute call is made on the library here."""
import ctypes
import os
import re

import numpy as np
import pytest

from helpers import ba_case_from_npz, load_golden, problem_from_case

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope='module')
def lib():
    import sfm_b200
    if not os.path.exists(sfm_b200.LIB_PATH):
        import __graft_entry__ as ge
        ge.build()
    return sfm_b200.load_library()


def _header_symbols():
    src = open(os.path.join(ROOT, 'include', 'sfmba.h')).read()
    return sorted(set(re.findall(r'SFMBA_API\s+[\w\s\*]+?\b(sfmba_\w+)\s*\(', src)))


def test_library_exports_every_declared_symbol(lib):
    from sfm_b200 import _capi
    names = _header_symbols()
    assert len(names) >= 19
    for n in names:
        assert hasattr(lib, n), 'libsfmba.so does not export %s' % n
    # the ctypes binding types exactly the declared set
    assert sorted(_capi.SYMBOLS) == names
    assert lib.sfmba_version() == 100


def test_struct_layouts_match_header(lib):
    from sfm_b200 import _capi
    # sizes implied by include/sfmba.h (LP64): 4 i32 + 7 pointers; options; result
    assert ctypes.sizeof(_capi.Problem) == 16 + 7 * 8
    opt = _capi.default_options()
    assert (opt.lambda_, opt.iterations, opt.mode, opt.dvdc_sign_quirk) == (5.0, 3, 0, 1)   # ba_processor.py:24
    # solver defaults (sfmba_default_options fills the struct on the C side: a wrong field offset would show here)
    assert (opt.solver, opt.precond, opt.pcg_tol, opt.pcg_max_iter) == (_capi.SOLVER_PCG, _capi.PRECOND_JACOBI, 1e-8, 0)
    assert (opt.compute_costs, opt.force_slow_path, opt.huber_delta) == (1, 0, 0.0)
    assert (opt.lm_lambda_min, opt.lm_lambda_max) == (1e-12, 1e12)
    assert ctypes.sizeof(_capi.Options) == 8 + 4 * 4 + 8 + 4 * 4 + 3 * 8 + 2 * 4
    assert ctypes.sizeof(_capi.Result) == 3 * 8 + 6 * 4 + 64 * 8 + 6 * 4 + 4 * 4


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip('a GPU is present')
    import sfm_b200
    with pytest.raises(sfm_b200.SfmbaError) as ei:
        sfm_b200.Engine(device=0)
    assert ei.value.code == -6 and 'no CPU fallback' in str(ei.value)
    tp = sfm_b200.TriangulationProcessor()
    z = load_golden('tri_inline.npz')
    with pytest.raises(sfm_b200.SfmbaError):
        tp.nonlinear_triangulate(z['linear'], [z['projs'][0], z['projs'][1]], [z['pairs'][0], z['pairs'][1]])


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, 'structure-from-motion_b200')
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(('.py', '.cu', '.cuh', '.h')):
                txt = open(os.path.join(dirpath, f)).read()
                assert 'sfm_oracle' not in txt and 'oracle/' not in txt, f


def test_host_visibility_index_is_bit_exact(oracle):
    import sfm_b200
    rng = np.random.default_rng(11)
    for trial in range(20):
        V = int(rng.integers(1, 7))
        N = int(rng.integers(1, 50))
        rows = []
        for v in range(V):
            n = int(rng.integers(0, 80))
            row = rng.integers(-1, N + 3, size=n)          # includes out-of-range ids
            row[rng.random(n) < 0.4] = -1
            rows.append(row)
        pt, cam, key, rowptr = sfm_b200.build_visibility_host(rows, N)
        lit = [(t, v, oracle.is_visible_literal(np.asarray(rows[v]), t)) for t in range(N) for v in range(V)]
        lit = np.array([x for x in lit if x[2] != -1], dtype=np.int64).reshape(-1, 3)
        assert np.array_equal(np.stack([pt, cam, key], 1).astype(np.int64), lit)
        assert rowptr[0] == 0 and rowptr[-1] == pt.shape[0] and np.all(np.diff(rowptr) >= 0)
        assert np.array_equal(np.repeat(np.arange(N), np.diff(rowptr)), pt)


@pytest.mark.parametrize('call', [1, 4])
def test_problem_packing_matches_recorded_call(oracle, call):
    case = ba_case_from_npz(load_golden('upenn_ba_call%d.npz' % call))
    prob, _, _ = problem_from_case(case)
    assert np.array_equal(np.stack([prob.obs_pt, prob.obs_cam, prob.obs_key], 1), case['vis'])
    assert prob.obs_uv.dtype == np.float32 and prob.cams.shape == (case['V'], 7)
    # the normalisation the device applies (ba_processor.py:339-342) reproduces the recorded b
    h = np.stack([prob.obs_uv[0], prob.obs_uv[1], np.ones(prob.num_obs)], 1).astype(np.float64)
    c = np.einsum('mab,mb->ma', prob.kinv.reshape(-1, 3, 3)[prob.obs_cam], h)
    assert np.max(np.abs(c[:, 0:2] / c[:, 2:3] - case['b0'])) < 1e-14
    assert np.max(np.abs(oracle.quaternion_to_rotation(prob.cams[:, 3:7]) - case['rot'])) < 1e-12


def test_utils_follow_reference_semantics(oracle):
    import sfm_b200
    from sfm_b200 import utils
    rng = np.random.default_rng(5)
    for _ in range(20):
        q = rng.normal(size=4)
        q[0] = abs(q[0]) + 0.2
        q /= np.linalg.norm(q)
        r = utils.convert_quaternion_to_rotation(q)
        assert np.allclose(r, oracle.quaternion_to_rotation(q), atol=0, rtol=0)
        assert np.allclose(utils.convert_rotation_to_quaternion(r).reshape(4), q, atol=1e-12)
    with pytest.raises(ValueError):
        utils.convert_rotation_to_quaternion(np.diag([1.0, -1.0, -1.0]))     # qw == 0 (utils.py:49-51)
    with pytest.raises(ValueError):
        utils.convert_rotation_to_quaternion(2 * np.eye(3))
    assert not utils.verify_rotation_mat(np.eye(3) * 1.001)
    assert sfm_b200.KeyPt(4).shape == (3, 4) and np.all(sfm_b200.KeyPt(4)[2] == 1)
    assert sfm_b200.TriPt(5).shape == (4, 5) and np.all(sfm_b200.TriPt(5)[3] == 1)
    import random
    random.seed(123)
    utils.RansacConfig(2.0, 0.99, 0.7, 6, 10)
    a = random.random()
    random.seed(-1)
    assert a == random.random()                                              # ctor reseeds with -1 (utils.py:172-174)


def test_falsy_kwargs_and_argument_errors():
    import sfm_b200
    tp = sfm_b200.TriangulationProcessor(damping_factor=0.25, iteration=7)
    with pytest.raises(ValueError):
        tp.triangulate([np.eye(3, 4)], [np.ones((3, 2)), np.ones((3, 2))])   # triangulation_processor.py:68
    with pytest.raises(ValueError):
        tp.triangulate([np.eye(3, 4)], [np.ones((3, 2))])                    # :73 fewer than 2 projections
    empty = tp.nonlinear_triangulate(np.zeros((4, 0)), [np.eye(3, 4)] * 2, [np.zeros((3, 0))] * 2, 0, 0)
    assert empty.shape == (4, 0)
    # size checks of linear_triangulate print and return None like the reference (triangulation_processor.py:122-134)
    assert tp.linear_triangulate([np.eye(3, 4)] * 2, [np.ones((3, 2)), np.ones((3, 3))]) is None
    cp = sfm_b200.CamposeProcessor(None, 5, 10)
    with pytest.raises(ValueError):
        cp.nonlinear_estimate_cam_pose_pnp(np.ones((3, 7)), np.ones((4, 8)), np.eye(3), np.eye(3), np.zeros((3, 1)))


def test_synthetic_shapes_and_sharding():
    from sfm_b200 import synthetic
    p1, t1 = synthetic.make_shape('ladybug')
    p2, _ = synthetic.make_shape('ladybug')
    V, N, m, _ = synthetic.SHAPES['ladybug']
    assert (p1.num_cams, p1.num_pts, p1.num_obs) == (V, N, m)
    assert np.array_equal(p1.obs_cam, p2.obs_cam) and np.array_equal(p1.pts, p2.pts)
    assert np.all(np.diff(p1.rowptr) >= 2) and np.unique(p1.obs_cam).size == V
    # point-major / view-minor order (ba_processor.py:304-306)
    key = p1.obs_pt.astype(np.int64) * V + p1.obs_cam
    assert np.all(np.diff(key) > 0)
    for world in (2, 3, 8):
        shards = [synthetic.shard_problem(p1, world, r) for r in range(world)]
        assert sum(s.num_pts for s in shards) == N and sum(s.num_obs for s in shards) == m
        assert np.array_equal(np.concatenate([s.obs_cam for s in shards]), p1.obs_cam)
        assert np.array_equal(np.concatenate([s.pts for s in shards]), p1.pts)
        assert max(s.num_obs for s in shards) - min(s.num_obs for s in shards) <= 2 * int(np.diff(p1.rowptr).max())
        for s in shards:
            assert s.rowptr[0] == 0 and s.rowptr[-1] == s.num_obs and s.num_cams == V


def test_lapack_sign_helper_matches_oracle(oracle):
    """CamposeProcessor._lapack_loc_signs (host, batched SVD) reads the same sign per hypothesis as the oracle's
    hypothesis-by-hypothesis restatement of campose_processor.py:616-631."""
    import sfm_b200
    from helpers import load_golden
    g = load_golden('ransac.npz')
    key, tri, k, samples = g['pnp_outliers_key'], g['pnp_tri'], g['pnp_k'], g['pnp_outliers_samples']
    ora = oracle.pnp_ransac(key, tri, k, samples[:40], 30.0)
    sg = sfm_b200.CamposeProcessor._lapack_loc_signs(key, tri, k, samples[:40])
    assert np.array_equal(sg, ora['signs'])


def test_epipolar_normalize_matches_oracle(oracle):
    import sfm_b200
    from helpers import load_golden
    g = load_golden('ransac.npz')
    left, right = g['fund_clean_left'], g['fund_clean_right']
    kl, kr = sfm_b200.KeyPt(left.shape[0]), sfm_b200.KeyPt(left.shape[0])
    kl[0:2, :] = left.T
    kr[0:2, :] = right.T
    pairs, tl, tr = sfm_b200.EpipolarProcessor._normalize([kl, kr])
    p2, tl2, tr2 = oracle.fund_normalize(left, right)
    assert np.array_equal(pairs, p2) and np.array_equal(tl, tl2) and np.array_equal(tr, tr2)
