"""Two-rank NCCL run of the point-sharded solve (needs 2 GPUs; skipped otherwise): every rank must end
with the cameras of the single-GPU solve and its own shard of the points."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, %(root)r)
import sfm_b200
from sfm_b200 import synthetic
rank, world = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
torch.cuda.set_device(rank)
dist.init_process_group('nccl', device_id=torch.device('cuda', rank))
buf = torch.zeros(128, dtype=torch.uint8, device='cuda')
if rank == 0:
    buf.copy_(torch.frombuffer(bytearray(sfm_b200.Engine.nccl_unique_id()), dtype=torch.uint8))
dist.broadcast(buf, 0)
eng = sfm_b200.Engine(device=rank, world_size=world, rank=rank, nccl_unique_id=bytes(buf.cpu().numpy().tobytes()))
full, _ = synthetic.make_shape('ladybug')
sh = synthetic.shard_problem(full, world, rank)
from sfm_b200 import _capi
res = eng.ba_solve_host(sh, lam=5.0, iterations=3, matvec_precision=_capi.MATVEC_F64)
sh2 = synthetic.shard_problem(full, world, rank)
res2 = eng.ba_solve_host(sh2, lam=5.0, iterations=3)              # default precision (AUTO)
sh3 = synthetic.shard_problem(full, world, rank)                  # adaptive LM: the cost of every candidate step and the
res3 = eng.ba_solve_host(sh3, lam=1e-2, iterations=6, mode=_capi.MODE_ADAPTIVE_LM, dvdc_sign_quirk=False,  # model decrease
                         matvec_precision=_capi.MATVEC_F64)       # are summed over the ranks
np.savez(os.path.join(%(out)r, 'rank%%d.npz' %% rank), cams=sh.cams, pts=sh.pts, cost=res.final_cost, c0=res.costs[0],
         cams_auto=sh2.cams, pts_auto=sh2.pts, cost_auto=res2.final_cost,
         cams_lm=sh3.cams, cost_lm=res3.final_cost, lam_lm=res3.final_lambda, acc_lm=res3.iterations_accepted)
eng.close()
dist.destroy_process_group()
'''


@pytest.mark.parametrize('peer', ['1', '0'])
def test_two_rank_solve_matches_single_gpu(engine, tmp_path, peer):
    """peer = '1': partial sums through the CUDA-IPC peer windows (flag-in-data words inside the solver's kernels);
    peer = '0' (SFMBA_PEER=0): the NCCL all-reduce fallback."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip('needs 2 GPUs')
    from sfm_b200 import synthetic
    script = tmp_path / 'worker.py'
    script.write_text(WORKER % dict(root=ROOT, out=str(tmp_path)))
    env = dict(os.environ, MASTER_ADDR='127.0.0.1', SFMBA_PEER=peer)
    subprocess.run([sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node=2',
                    '--master-addr', '127.0.0.1', '--master-port', '29533', str(script)], check=True, env=env,
                   timeout=600)
    full, _ = synthetic.make_shape('ladybug')
    from sfm_b200 import _capi
    full_auto, _ = synthetic.make_shape('ladybug')
    res_auto = engine.ba_solve_host(full_auto, lam=5.0, iterations=3)
    res = engine.ba_solve_host(full, lam=5.0, iterations=3, matvec_precision=_capi.MATVEC_F64)
    out = [np.load(tmp_path / ('rank%d.npz' % r)) for r in range(2)]
    for o in out:
        assert np.max(np.abs(o['cams'] - full.cams)) < 1e-9
        assert abs(float(o['cost']) - res.final_cost) <= 1e-9 * res.final_cost
        assert abs(float(o['c0']) - res.costs[0]) <= 1e-9 * res.costs[0]
    assert np.max(np.abs(np.concatenate([o['pts'] for o in out]) - full.pts)) < 1e-8
    # default precision: FP32 sums over another partition of the points agree to FP32 accuracy
    for o in out:
        assert np.max(np.abs(o['cams_auto'] - full_auto.cams)) < 1e-6
        assert abs(float(o['cost_auto']) - res_auto.final_cost) <= 1e-6 * res_auto.final_cost
    assert np.max(np.abs(np.concatenate([o['pts_auto'] for o in out]) - full_auto.pts)) < 1e-6
    # adaptive LM: same accept / reject decisions, same damping, same answer
    full_lm, _ = synthetic.make_shape('ladybug')
    res_lm = engine.ba_solve_host(full_lm, lam=1e-2, iterations=6, mode=_capi.MODE_ADAPTIVE_LM, dvdc_sign_quirk=False,
                                  matvec_precision=_capi.MATVEC_F64)
    for o in out:
        assert int(o['acc_lm']) == res_lm.iterations_accepted
        assert abs(float(o['lam_lm']) - res_lm.final_lambda) <= 1e-6 * res_lm.final_lambda
        assert abs(float(o['cost_lm']) - res_lm.final_cost) <= 1e-8 * res_lm.final_cost
        assert np.max(np.abs(o['cams_lm'] - full_lm.cams)) < 1e-7


TIMEOUT_WORKER = r'''
import os, sys
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, %(root)r)
import sfm_b200
from sfm_b200 import _capi, synthetic
rank, world = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
torch.cuda.set_device(rank)
dist.init_process_group('nccl', device_id=torch.device('cuda', rank))
buf = torch.zeros(128, dtype=torch.uint8, device='cuda')
if rank == 0:
    buf.copy_(torch.frombuffer(bytearray(sfm_b200.Engine.nccl_unique_id()), dtype=torch.uint8))
dist.broadcast(buf, 0)
eng = sfm_b200.Engine(device=rank, world_size=world, rank=rank, nccl_unique_id=bytes(buf.cpu().numpy().tobytes()))
full, _ = synthetic.make_shape('small')
sh = synthetic.shard_problem(full, world, rank)
eng.ba_solve_host(sh, lam=5.0, iterations=1)            # both ranks: the peer windows get mapped
dist.barrier()
out = 'skipped'
if rank == 0:                                           # rank 1 never enters the second solve
    before = sh.cams.copy()
    try:
        # (no final cost: that one number is all-reduced by NCCL, which has its own way of dealing with a dead peer)
        eng.ba_solve_host(sh, lam=5.0, iterations=2, compute_costs=False)
        out = 'returned OK'
    except _capi.SfmbaError as e:
        out = 'code %%d unchanged %%s' %% (e.code, bool(np.array_equal(before, sh.cams)))
open(os.path.join(%(out)r, 'timeout%%d.txt' %% rank), 'w').write(out)
dist.barrier()
os._exit(0)                                             # the windows are out of step now: no orderly teardown
'''


def test_missing_peer_is_reported_as_comm_error(engine, tmp_path):
    """A rank whose peer never arrives at an exchange gives up after the timeout, returns SFMBA_ERR_COMM (not SFMBA_OK
    with corrupted sums) and leaves the cameras of the failed iteration untouched."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip('needs 2 GPUs')
    from sfm_b200 import _capi
    script = tmp_path / 'worker_timeout.py'
    script.write_text(TIMEOUT_WORKER % dict(root=ROOT, out=str(tmp_path)))
    env = dict(os.environ, MASTER_ADDR='127.0.0.1', SFMBA_PEER_TIMEOUT_MS='1500')
    subprocess.run([sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node=2',
                    '--master-addr', '127.0.0.1', '--master-port', '29534', str(script)], check=True, env=env,
                   timeout=300)
    got = (tmp_path / 'timeout0.txt').read_text()
    assert got == 'code %d unchanged True' % _capi.ERR_COMM, got
