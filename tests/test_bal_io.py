"""BAL text reader / writer and the conversion to the reference's camera model."""
import numpy as np
import pytest


def _bal():
    from sfm_b200 import bal
    return bal


def test_bal_write_read_roundtrip(tmp_path):
    bal = _bal()
    b = bal.synthetic_bal(num_cams=5, num_pts=60, seed=3)
    p = tmp_path / 'problem.txt'
    bal.write_bal(p, b)
    r = bal.read_bal(p)
    assert np.array_equal(r.cams, b.cams) and np.array_equal(r.pts, b.pts)
    assert np.array_equal(r.obs_cam, b.obs_cam) and np.array_equal(r.obs_pt, b.obs_pt)
    assert np.array_equal(r.obs_xy, b.obs_xy)
    with open(p, 'a') as f:
        f.write('1.0\n')
    with pytest.raises(ValueError):
        bal.read_bal(p)


def test_bal_empty_and_ragged(tmp_path):
    bal = _bal()
    b = bal.BalFile(cams=np.zeros((2, 9)), pts=np.zeros((3, 3)), obs_cam=np.zeros(0, dtype=np.int32),
                    obs_pt=np.zeros(0, dtype=np.int32), obs_xy=np.zeros((0, 2)))
    b.cams[:, 6] = 500.0
    p = tmp_path / 'empty.txt'
    bal.write_bal(p, b)
    r = bal.read_bal(p)
    assert r.num_obs == 0 and r.num_cams == 2 and r.num_pts == 3
    prob, order = bal.bal_to_problem(r)
    assert prob.num_obs == 0 and list(prob.rowptr) == [0, 0, 0, 0]


def test_bal_conversion_matches_bal_model(oracle):
    """Exact BAL observations (with radial distortion) have zero residual in the converted problem, and a noisy
    file's cost in normalised coordinates is the undistorted pixel residual over f."""
    bal = _bal()
    b = bal.synthetic_bal(num_cams=7, num_pts=200, seed=1, noise_px=0.0)
    prob, order = bal.bal_to_problem(b)
    assert np.all(np.diff(prob.obs_pt) >= 0)                                  # point-major
    same = np.diff(prob.obs_pt) == 0
    assert np.all(np.diff(prob.obs_cam)[same] > 0)                            # view-minor
    obs_b = oracle.normalise_observations(prob.obs_uv.T.astype(np.float64), prob.obs_cam,
                                           np.linalg.inv(prob.kinv.reshape(-1, 3, 3)))
    cost = oracle.ba_cost(prob.cams, prob.pts, prob.obs_cam, prob.obs_pt, obs_b)
    assert cost < 1e-9 * prob.num_obs                                         # float32 pixel rounding only
    # the quaternions reproduce the rotations, including cameras turned by half a turn
    R = bal.rodrigues_to_rot(np.array([[0, 0, np.pi], [np.pi, 0, 0], [0.1, 3.0, -0.2], [0, 0, 0], [1e-14, 0, 0]]))
    q = bal.rot_to_quat(R)
    assert np.max(np.abs(bal._quat_to_rot(q) - R)) < 1e-12
    for i in range(R.shape[0]):
        assert np.max(np.abs(oracle.quaternion_to_rotation(q[i]) - R[i])) < 1e-12
    assert np.max(np.abs(bal.rodrigues_to_rot(bal.rot_to_rodrigues(R)) - R)) < 1e-9


def test_bal_problem_roundtrip():
    bal = _bal()
    b = bal.synthetic_bal(num_cams=6, num_pts=100, seed=2)
    prob, _ = bal.bal_to_problem(b)
    back = bal.problem_to_bal(prob, b)
    assert np.max(np.abs(back.cams - b.cams)) < 1e-10
    assert np.max(np.abs(back.pts - b.pts)) < 1e-12
    # undistortion inverts the BAL distortion
    rng = np.random.default_rng(0)
    p = rng.uniform(-0.6, 0.6, (1000, 2))
    k1, k2 = rng.normal(0, 0.05, 1000), rng.normal(0, 0.01, 1000)
    assert np.max(np.abs(bal.undistort(bal.distort(p, k1, k2), k1, k2) - p)) < 1e-13


@pytest.mark.gpu
def test_bal_file_solve_on_gpu(engine, tmp_path):
    """A perturbed BAL file, read from disk, solved with the adaptive LM (exact Jacobian), written back: the
    pixel residual of the BAL model falls to the noise level."""
    from sfm_b200 import _capi
    bal = _bal()
    b = bal.synthetic_bal(num_cams=10, num_pts=800, seed=4, noise_px=0.3)
    rng = np.random.default_rng(9)
    b.pts = b.pts + rng.normal(0, 0.05, b.pts.shape)
    b.cams[:, 3:6] += rng.normal(0, 0.02, (b.num_cams, 3))
    b.cams[:, 0:3] += rng.normal(0, 0.003, (b.num_cams, 3))
    p = tmp_path / 'noisy.txt'
    bal.write_bal(p, b)
    r = bal.read_bal(p)
    rms0 = np.sqrt(np.mean(np.sum((bal.bal_project(r) - r.obs_xy) ** 2, axis=1)))
    prob, _ = bal.bal_to_problem(r)
    res = engine.ba_solve_host(prob, lam=1e-3, iterations=30, mode=_capi.MODE_ADAPTIVE_LM, dvdc_sign_quirk=False)
    out = bal.problem_to_bal(prob, r)
    rms1 = np.sqrt(np.mean(np.sum((bal.bal_project(out) - out.obs_xy) ** 2, axis=1)))
    assert res.final_cost < 0.05 * res.initial_cost
    assert rms0 > 5.0 and rms1 < 0.6, (rms0, rms1)
    bal.write_bal(tmp_path / 'solved.txt', out)
    assert np.array_equal(bal.read_bal(tmp_path / 'solved.txt').cams, out.cams)
