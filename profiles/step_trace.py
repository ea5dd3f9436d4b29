#!/usr/bin/env python3
"""Where the time of one k_pcg_step launch goes (profiling build only).

    python profiles/build_variant.py trace ba_pcg.cu -DSFMBA_STEP_TRACE
    SFMBA_LIB=variants/libsfmba_trace.so python profiles/step_trace.py [shape]

Thread 0 of every CTA stamps globaltimer and clock64 at the phase boundaries of k_pcg_step (STEP_STAMP in
csrc/ba_pcg.cu); the stamps of the LAST launch of the solve are read back.  Printed: per phase, the median / max over
CTAs of the clock64 difference (us at the SM clock), and the span of the launch on the global timer."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sfm_b200                      # noqa: E402
from sfm_b200 import synthetic       # noqa: E402

NAMES = ['entry', 'prefetch issued', 'pdl_wait', 'gather', 'exchange (mode 3), w, dots', 'block sums', 'grid barrier',
         'partials + alpha', 'update']

shape = sys.argv[1] if len(sys.argv) > 1 else 'venice'
world = int(os.environ.get('WORLD_SIZE', '1'))
rank = int(os.environ.get('RANK', '0'))
local = int(os.environ.get('LOCAL_RANK', '0'))
nccl_id = None
if world > 1:                        # under torchrun: points sharded over the ranks, the exchange inside the kernel
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    b = torch.zeros(128, dtype=torch.uint8, device='cuda')
    if rank == 0:
        b.copy_(torch.frombuffer(bytearray(sfm_b200.Engine.nccl_unique_id()), dtype=torch.uint8))
    dist.broadcast(b, 0)
    nccl_id = bytes(b.cpu().numpy().tobytes())
eng = sfm_b200.Engine(device=local, world_size=world, rank=rank, nccl_unique_id=nccl_id)
prob, _ = synthetic.make_shape(shape)
if world > 1:
    prob = synthetic.shard_problem(prob, world, rank)
t = eng.to_device(prob)
c0, p0 = t['cams'].clone(), t['pts'].clone()
for _ in range(3):
    t['cams'].copy_(c0)
    t['pts'].copy_(p0)
    res = eng.ba_solve_device(t, lam=5.0, iterations=3, compute_costs=False)
if rank != 0:                        # the other ranks report into their own file (TRACE_ALL_RANKS=dir)
    if not os.environ.get('TRACE_ALL_RANKS'):
        eng.close()
        sys.exit(0)
    sys.stdout = open(os.path.join(os.environ['TRACE_ALL_RANKS'], 'step_trace_rank%d.txt' % rank), 'w')
buf = np.zeros((2, 256, 16), dtype=np.uint64)
rc = eng.lib.sfmba_debug_step_trace(buf.ctypes.data_as(C.c_void_p))
assert rc == 0, rc
mhz = 1965.0
g, c = buf[0].astype(np.int64), buf[1].astype(np.int64)
n = int((g[:, 0] > 0).sum())
g, c = g[:n], c[:n]
print('%d CTAs, %d PCG iterations in the solve' % (n, res.pcg_iterations))
if os.environ.get('TRACE_ALL_RANKS'):
    np.save(os.path.join(os.environ['TRACE_ALL_RANKS'], 'step_trace_raw_rank%d.npy' % rank), buf)
t_first = g[:, 0].min()
print('launch span on the global timer: first entry -> last stamp %.2f us' % ((g[:, :9].max() - t_first) / 1e3))
for i in range(1, 9):
    dt = (c[:, i] - c[:, i - 1]) / mhz
    at = (g[:, i] - t_first) / 1e3
    print('%-22s  dt median %6.2f  max %6.2f us   | reached at (global) median %6.2f  max %6.2f us' %
          (NAMES[i], np.median(dt), dt.max(), np.median(at), at.max()))
if c[:, 13].max() > 0:
    first = (c[:, 13] - c[:, 2]) / mhz
    total = (c[:, 3] - c[:, 2]) / mhz
    rows = c[:, 12]
    print('gather of warp 0: rows median %d (min %d, max %d); first batch done after %.2f us (median), whole gather %.2f us' %
          (np.median(rows), rows.min(), rows.max(), np.median(first), np.median(total)))
    print('  lane sums combined (before y = f(sums)) at %.2f us after the wait (median)' % np.median((c[:, 14] - c[:, 2]) / mhz))
    A = np.vstack([rows, np.ones_like(rows)]).T.astype(float)
    k, b = np.linalg.lstsq(A, total, rcond=None)[0]
    print('  fit: %.3f us + %.4f us per row (%.2f us per 160-row batch)' % (b, k, 160 * k))
print('scalars written by CTA 0 %.2f us after its update' % ((c[0, 11] - c[0, 8]) / mhz))

# ---- the fused camera finish (k_cam_finish), same build
if hasattr(eng.lib, 'sfmba_debug_fin_trace') or True:
    try:
        fb = np.zeros((2, 512, 8), dtype=np.uint64)
        eng2 = None
        rc = C.CDLL(sfm_b200._capi.LIB_PATH).sfmba_debug_fin_trace(fb.ctypes.data_as(C.c_void_p))
        fg, fc = fb[0].astype(np.int64), fb[1].astype(np.int64)
        nf = int((fg[:, 0] > 0).sum())
        fg, fc = fg[:nf], fc[:nf]
        names = ['entry', 'cost sum / scalars', 'row gather', 'Cholesky', 'inverse columns', 'rhs, rows, stores', 'block_sum + ticket']
        t0 = fg[:, 0].min()
        print('k_cam_finish: %d CTAs, span %.2f us' % (nf, (fg[:, :7].max() - t0) / 1e3))
        for i in range(1, 7):
            dt = (fc[:, i] - fc[:, i - 1]) / mhz
            at = (fg[:, i] - t0) / 1e3
            print('%-22s  dt median %6.2f  max %6.2f us   | reached at (global) median %6.2f  max %6.2f us' %
                  (names[i], np.median(dt), dt.max(), np.median(at), at.max()))
    except AttributeError:
        pass
eng.close()
