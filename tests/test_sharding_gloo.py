"""The N > 1 decomposition on CPU: world_size-2 `gloo` processes each take the point shard
`synthetic.shard_problem` gives them, form their share of the reduced camera system with the oracle,
all-reduce exactly the quantities libsfmba all-reduces over NCCL (camera blocks A_i / e_p / the Schur
sums, then the cost), and must reproduce the unsharded step."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, lam, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, 'oracle'))
    import torch
    import torch.distributed as dist
    import sfm_oracle as oracle
    from sfm_b200 import synthetic

    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    full, _ = synthetic.make_shape('tiny')
    sh = synthetic.shard_problem(full, world, rank)
    V = full.num_cams
    b = sh.obs_uv.T.astype(np.float64)
    oc, op = sh.obs_cam.astype(np.int64), sh.obs_pt.astype(np.int64)
    ne = oracle.ba_normal_equations(sh.cams, sh.pts, oc, op, b, lam)
    # local share of S and g: the damping on A is added once (after the reduction), like libsfmba does
    S_loc, g_loc = oracle._reduced_system_dense(ne, oc, op, V)
    S_loc -= lam * np.eye(7 * V)
    cost_loc = float(np.sum(np.square(ne['r'])))
    msg = torch.from_numpy(np.concatenate([S_loc.reshape(-1), g_loc, [cost_loc]]))
    dist.all_reduce(msg)
    msg = msg.numpy()
    S = msg[:49 * V * V].reshape(7 * V, 7 * V) + lam * np.eye(7 * V)
    g = msg[49 * V * V:49 * V * V + 7 * V]
    cost = msg[-1]
    dp = np.linalg.solve(S, g).reshape(V, 7)
    bt_dp = oracle._segment_sum(np.einsum('mij,mi->mj', ne['W'], dp[oc]), op, sh.num_pts)
    dx = np.einsum('nij,nj->ni', ne['Dinv'], ne['e_x'] - bt_dp)
    cams = sh.cams + dp
    cams[:, 3:7] /= np.sqrt(np.sum(np.square(cams[:, 3:7]), axis=1, keepdims=True))
    pts = sh.pts + dx
    np.savez(os.path.join(out_dir, 'rank%d.npz' % rank), cams=cams, pts=pts, cost=cost)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_point_sharded_step_equals_unsharded_step(oracle, tmp_path):
    import torch.multiprocessing as mp
    from sfm_b200 import synthetic
    world, lam = 2, 5.0
    port = _free_port()
    mp.spawn(_worker, args=(world, port, lam, str(tmp_path)), nprocs=world, join=True)
    full, _ = synthetic.make_shape('tiny')
    b = full.obs_uv.T.astype(np.float64)
    cams, pts, costs = oracle.ba_solve(full.cams, full.pts, full.obs_cam, full.obs_pt, b, lam, 1)
    out = [np.load(os.path.join(str(tmp_path), 'rank%d.npz' % r)) for r in range(world)]
    for o in out:                                   # cameras are replicated: every rank holds the same update
        assert np.max(np.abs(o['cams'] - cams)) < 1e-10
        assert abs(float(o['cost']) - costs[0]) < 1e-10 * costs[0]
    assert np.max(np.abs(np.concatenate([o['pts'] for o in out]) - pts)) < 1e-9
