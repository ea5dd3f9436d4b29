#!/usr/bin/env python3
"""bench.py -- BA LM iterations/s at Venice shape (BASELINE.json), 1..8 B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--shape venice]

A "step" is one BaProcessor-style BA call in reference mode (lambda 5, 3 always-accepted LM
iterations, reduced system solved to pcg_tol) on the synthetic Venice-shaped problem
(1778 cameras / 993 923 points / 5 001 946 observations); every step starts from the same initial
parameters (device-to-device restore inside the timed region).
  value   whole-job LM iterations/s with the problem resident in HBM
  e2e     the same through sfmba_ba_solve_host (HOST buffers: pinned H2D of every input, D2H of
          cameras and points inside the timed region)
  roofline      the implicit-Schur matvec K3 (one PCG iteration of k_pcg_solve), algorithmic bytes / CUDA-event time
  cpu_baseline  the NumPy/SciPy oracle port timed on this host on a 1/4-scale Venice sample
With N > 1 (torchrun) the points are sharded by observation count across ranks (strong scaling),
cameras replicated, NCCL all-reduces the camera-side blocks and the matvec result.
`--impl reference` times the oracle port (the reference is pure Python and cannot travel; its
literal code cannot run beyond ~3 k observations) on the host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

LM_ITERS = 3
LAMBDA = 5.0
CPU_SAMPLE_SCALE = 0.25
# ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum of ONE k_matvec<0> launch at the full Venice
# shape on one GPU (profiles/r1_ncu_full_v55_summary.txt, mean of two launches: 104.52 MB read + 19.65 MB written); the same for the one
# k_linearize<0, 0> launch of that capture (155.09 MB read + 133.70 MB written)
NCU_DRAM_BYTES_PER_MATVEC = {('venice', 1.0, 1): 124.17e6}
NCU_DRAM_BYTES_PER_LINEARIZE = {('venice', 1.0, 1): 288.79e6}


def _peaks():
    try:
        with open(os.path.join(ROOT, 'MEASURED_PEAKS.json')) as f:
            return float(json.load(f)['hbm_gbs']), 'measured'
    except Exception:
        return 6650.0, 'fallback'


class ClockSampler(threading.Thread):
    def __init__(self, index=0):
        super().__init__(daemon=True)
        self.index = index
        self.samples = []
        self.stop_flag = False

    def run(self):
        q = 'clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,' \
            'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap'
        while not self.stop_flag:
            try:
                out = subprocess.run(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + q,
                                      '--format=csv,noheader,nounits'], capture_output=True, text=True, timeout=5).stdout
                parts = [s.strip() for s in out.strip().split(',')]
                if len(parts) >= 6:
                    self.samples.append(parts)
            except Exception:
                pass
            time.sleep(0.02)

    def summary(self):
        if not self.samples:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': []}
        sm = sorted(float(s[0]) for s in self.samples)
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        reasons = [n for i, n in enumerate(names) if any(s[2 + i].lower().startswith('active') for s in self.samples)]
        return {'sm_mhz': sm[len(sm) // 2], 'sm_max_mhz': float(self.samples[0][1]), 'reasons': reasons,
                'samples': len(sm)}


def cpu_baseline(steps=1, warmup=0, shape='venice'):
    """Oracle port on a bounded sample of the workload, all host threads NumPy/BLAS will use."""
    sys.path.insert(0, os.path.join(ROOT, 'oracle'))
    import sfm_oracle as oracle
    from sfm_b200 import synthetic
    prob, _ = synthetic.make_shape(shape, scale=CPU_SAMPLE_SCALE)
    b = prob.obs_uv.T.astype(np.float64)
    V, N, m, _ = synthetic.SHAPES[shape]
    times = []
    for s in range(warmup + steps):
        t0 = time.perf_counter()
        oracle.ba_solve(prob.cams, prob.pts, prob.obs_cam, prob.obs_pt, b, LAMBDA, 1, sparse=True, return_costs=False)
        dt = time.perf_counter() - t0
        if s >= warmup:
            times.append(dt)
    per_iter = float(np.mean(times))
    obs_per_s = prob.num_obs / per_iter
    return {
        'value': obs_per_s / m,                 # LM iterations/s at full shape, linear extrapolation
        'unit': 'LM iterations/s',
        'cores': os.cpu_count(),
        'kind': 'port',
        'sample': '%d LM iteration(s) of oracle/sfm_oracle.py (NumPy + scipy.sparse, dense reduced solve) on a '
                  '%.2f-scale %s problem (%d cams / %d pts / %d obs, %.1f s each); obs/s scaled linearly to the '
                  'full shape, which favours the CPU (its reduced solve is cubic in cameras)'
                  % (steps, CPU_SAMPLE_SCALE, shape, prob.num_cams, prob.num_pts, prob.num_obs, per_iter),
        'obs_per_s': obs_per_s,
        'ms_per_step': per_iter * 1e3,
    }


def workload_name(shape, V, N, m, pcg_tol):
    return ('%s-shaped BA: %d cams / %d pts / %d obs; reference mode: lambda=%g, %d LM iterations per step, PCG to %.0e'
            % (shape, V, N, m, LAMBDA, LM_ITERS, pcg_tol))


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    base = cpu_baseline(steps=max(1, args.steps), warmup=args.warmup, shape=args.shape)
    from sfm_b200 import synthetic
    V, N, m, _ = synthetic.SHAPES[args.shape]
    line = {
        'impl': 'reference', 'metric': 'ba_lm_iterations_per_sec', 'value': base['value'], 'unit': 'LM iterations/s',
        'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': base['ms_per_step'],
        'higher_is_better': True, 'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
        'config': {'workload': workload_name(args.shape, V, N, m, args.pcg_tol),
                   'note': 'CPU arm: each timed step is ONE LM iteration of the oracle port (exact reduced solve instead '
                           'of PCG) on a %.2f-scale sample of the workload; value = LM iterations/s scaled linearly in '
                           'observations to the full shape' % CPU_SAMPLE_SCALE,
                   'sample_scale': CPU_SAMPLE_SCALE},
        'cpu_baseline': {k: base[k] for k in ('value', 'unit', 'cores', 'kind', 'sample')},
        'e2e': {'value': base['value'], 'unit': 'LM iterations/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'jacobian_schur_obs_per_sec': base['obs_per_s'],
    }
    emit(line)


def run_ours(args):
    import torch
    import torch.distributed as dist
    import sfm_b200
    from sfm_b200 import synthetic

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    nccl_id = None
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
        buf = torch.zeros(128, dtype=torch.uint8, device='cuda')
        if rank == 0:
            buf.copy_(torch.frombuffer(bytearray(sfm_b200.Engine.nccl_unique_id()), dtype=torch.uint8))
        dist.broadcast(buf, 0)
        nccl_id = bytes(buf.cpu().numpy().tobytes())
    eng = sfm_b200.Engine(device=local, world_size=world, rank=rank, nccl_unique_id=nccl_id)

    full, _ = synthetic.make_shape(args.shape, scale=args.scale)
    V, N, m = full.num_cams, full.num_pts, full.num_obs
    prob = synthetic.shard_problem(full, world, rank) if world > 1 else full
    opt = eng.make_options(lam=LAMBDA, iterations=LM_ITERS, pcg_tol=args.pcg_tol, compute_costs=True,
                           pcg_max_iter=args.pcg_max_iter,
                           matvec_precision={'auto': 0, 'f64': 1, 'f32': 2}[args.matvec])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident path ------------------------------------------------------------------
    t = eng.to_device(prob)
    cams0, pts0 = t['cams'].clone(), t['pts'].clone()
    # L2 flush between steps is unnecessary: one step streams > 1 GB per matvec, far beyond the 126 MB L2
    eng.profile_enable(args.profile_timed)
    res = None
    for _ in range(args.warmup):
        t['cams'].copy_(cams0)
        t['pts'].copy_(pts0)
        res = eng.ba_solve_device(t, opt)
    eng.profile_reset()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches = 0
    pcg_iters = 0
    phase = {'linearize': 0.0, 'camreduce': 0.0, 'pcg': 0.0, 'update': 0.0, 'index': 0.0}
    ev0.record()
    for _ in range(args.steps):
        t['cams'].copy_(cams0)
        t['pts'].copy_(pts0)
        res = eng.ba_solve_device(t, opt)
        launches += res.kernel_launches + 2
        pcg_iters += res.pcg_iterations
        for k in phase:
            phase[k] += res.ms[k]
    ev1.record()
    barrier()
    ms = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device='cuda')
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    if not args.profile_timed:
        # per-kernel CUDA-event timing adds event records between launches: one extra, untimed step
        eng.profile_enable(True)
        eng.profile_reset()
        t['cams'].copy_(cams0)
        t['pts'].copy_(pts0)
        res_prof = eng.ba_solve_device(t, opt)
        torch.cuda.synchronize()
        prof_pcg_iters = res_prof.pcg_iterations
    else:
        prof_pcg_iters = pcg_iters
    prof = {k: eng.profile_get(k) for k in ('linearize', 'matvec', 'pcg_tail', 'backsub', 'cost')}
    eng.profile_enable(False)

    # ---- end-to-end path: HOST (pinned) buffers through sfmba_ba_solve_host ----------------------
    def pinned_like(a):
        tt = torch.empty(a.shape, dtype=torch.from_numpy(a).dtype, pin_memory=True)
        tt.numpy()[...] = a
        return tt

    pin = {k: pinned_like(getattr(prob, k)) for k in ('cams', 'pts', 'obs_uv', 'obs_cam', 'obs_pt', 'rowptr')}
    hp = sfm_b200.BAProblem(cams=pin['cams'].numpy(), pts=pin['pts'].numpy(), obs_uv=pin['obs_uv'].numpy(),
                            obs_cam=pin['obs_cam'].numpy(), obs_pt=pin['obs_pt'].numpy(), rowptr=pin['rowptr'].numpy())
    assert hp.cams.ctypes.data == pin['cams'].data_ptr()
    cams_h0, pts_h0 = prob.cams.copy(), prob.pts.copy()
    h2d = sum(getattr(hp, k).nbytes for k in ('cams', 'pts', 'obs_uv', 'obs_cam', 'rowptr'))   # obs_pt is not uploaded
    d2h = hp.cams.nbytes + hp.pts.nbytes
    e2e_steps = max(1, min(args.steps, 5))
    for _ in range(min(args.warmup, 2)):
        hp.cams[...] = cams_h0
        hp.pts[...] = pts_h0
        eng.ba_solve_host(hp, opt)
    barrier()
    e2e_sum = 0.0
    for _ in range(e2e_steps):
        hp.cams[...] = cams_h0                   # harness: restore the inputs (the solve is in place); not timed
        hp.pts[...] = pts_h0
        barrier()
        t0 = time.perf_counter()
        r2 = eng.ba_solve_host(hp, opt)          # H2D of every input, solve, D2H of cameras and points
        e2e_sum += time.perf_counter() - t0
        launches_e2e = r2.kernel_launches
    barrier()
    e2e_s = torch.tensor([e2e_sum], dtype=torch.float64, device='cuda')
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_s = float(e2e_s.item())
    if rank == 0:
        sampler.stop_flag = True
        sampler.join(timeout=2)

    if rank == 0:
        peak, peak_kind = _peaks()
        lm_per_s = LM_ITERS * args.steps / (ms_total * 1e-3)
        m_local = prob.num_obs
        n_local = prob.num_pts
        k3_bytes = 84.0 * m_local + 28.0 * n_local + 112.0 * V       # SURVEY.md section 8(d), per PCG iteration
        pcg_per_lm = pcg_iters / max(1, LM_ITERS * args.steps)
        # The host queues PCG iterations ahead of the convergence flag, so a few launches per solve exit at
        # their first instruction.  Their (small) time is kept and the total is divided by the iterations
        # that did work: time per real PCG iteration, not an average diluted by the no-op launches.
        mv_a_ms, mv_a_n = prof['matvec']
        mv_b_ms, mv_b_n = prof['pcg_tail']
        k3_ms = (mv_a_ms + mv_b_ms) / max(prof_pcg_iters, 1)
        k3_name = 'K3 implicit-Schur matvec = k_matvec + k_pcg_tail (one PCG iteration, rank 0 shard)'
        k3_gbs = k3_bytes / (k3_ms * 1e-3) / 1e9 if k3_ms > 0 else 0.0
        lin_ms, lin_n = prof['linearize']
        cr_ms, cr_n = 0.0, 1
        k12_bytes = (104.0 + 92.0) * m_local + 48.0 * n_local + 308.0 * V
        k12_ms = (lin_ms / max(lin_n, 1)) + (cr_ms / max(cr_n, 1))
        k12_gbs = k12_bytes / (k12_ms * 1e-3) / 1e9 if k12_ms > 0 else 0.0
        line = {
            'metric': 'ba_lm_iterations_per_sec', 'value': lm_per_s, 'unit': 'LM iterations/s', 'n_gpus': world,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': ms_total / args.steps,
            'higher_is_better': True, 'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f64',
            'data': 'synthetic',
            'config': {'workload': workload_name(args.shape, V, N, m, args.pcg_tol),
                       'parallelism': 'points sharded over %d GPU(s), cameras replicated' % world,
                       'pcg_iterations_per_lm_iteration': pcg_iters / max(1, LM_ITERS * args.steps),
                       'preconditioner': 'block-Jacobi on A + lambda I (sfmba_options.precond default)',
                       'l2': 'inputs larger than L2 (>1 GB streamed per matvec)'},
            'jacobian_schur_obs_per_sec': (m / (k12_ms * 1e-3)) if k12_ms > 0 else None,
            'final_cost': res.final_cost, 'initial_cost': res.initial_cost,
            'phase_ms_per_step': {k: v / args.steps for k, v in phase.items()},
            'kernel_ms': {k: (v[0] / max(prof_pcg_iters if k in ('matvec', 'pcg_tail') else v[1], 1))
                          for k, v in prof.items() if v[1] > 0},
            'roofline': {'kernel': k3_name,
                         'bound': 'hbm', 'achieved': k3_gbs, 'peak': peak, 'unit': 'GB/s', 'frac': k3_gbs / peak,
                         'peak_kind': peak_kind, 'traffic': NCU_DRAM_BYTES_PER_MATVEC.get((args.shape, args.scale, world)),
                         'traffic_source': 'profiles/r1_ncu_full_v55_summary.txt (k_matvec<0, 0>, one launch)',
                         'algorithmic_bytes': k3_bytes, 'ms': k3_ms},
            'roofline_jacobian_schur': {'kernel': 'K1+K2 = k_linearize (residuals, Jacobians, point Schur blocks and camera Gram rows in one launch; J never stored)', 'bound': 'hbm',
                                        'achieved': k12_gbs, 'peak': peak, 'unit': 'GB/s', 'frac': k12_gbs / peak,
                                        'algorithmic_bytes': k12_bytes, 'ms': k12_ms,
                                        'traffic': NCU_DRAM_BYTES_PER_LINEARIZE.get((args.shape, args.scale, world)),
                                        'traffic_source': 'profiles/r1_ncu_full_v55_summary.txt (k_linearize<0, 0>, one launch)'},
            'e2e': {'value': LM_ITERS * e2e_steps / e2e_s, 'unit': 'LM iterations/s', 'h2d_bytes_per_step': h2d,
                    'd2h_bytes_per_step': d2h, 'steps': e2e_steps, 'ms_per_step': e2e_s / e2e_steps * 1e3,
                    'kernel_launches_per_step': launches_e2e},
            'gpu_launches': launches,
            'clocks': sampler.summary(),
        }
        if world == 1 and not args.no_cpu:
            base = cpu_baseline(steps=1, warmup=0, shape=args.shape)
            line['cpu_baseline'] = {k: base[k] for k in ('value', 'unit', 'cores', 'kind', 'sample')}
        emit(line)
    eng.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    # Libraries (NCCL_DEBUG=VERSION, torchrun banners) may print to stdout; the contract is ONE JSON line there.
    # Everything else goes to stderr: fd 1 is pointed at stderr and the JSON line is written to the saved fd.
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    def emit(obj):
        os.write(real_stdout, (json.dumps(obj) + '\n').encode())
    globals()['emit'] = emit
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--shape', default='venice')
    ap.add_argument('--scale', type=float, default=1.0)
    ap.add_argument('--pcg-tol', type=float, default=1e-8)
    ap.add_argument('--no-cpu', action='store_true')
    ap.add_argument('--matvec', default='auto', choices=['auto', 'f64', 'f32'],
                    help='sfmba_options.matvec_precision (auto = the library default)')
    ap.add_argument('--pcg-max-iter', type=int, default=0, help='diagnostic: cap PCG iterations (0 = library default)')
    ap.add_argument('--profile-timed', action='store_true', help='keep per-kernel event timing on inside the timed region')
    args = ap.parse_args()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
