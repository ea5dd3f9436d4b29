#!/usr/bin/env python3
"""bench.py -- BA LM iterations/s at Venice shape (BASELINE.json), 1..8 B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--shape venice] [--matvec auto|f64|f32]

A "step" is one BaProcessor-style BA call in reference mode (lambda 5, 3 always-accepted LM iterations, reduced system
solved to pcg_tol) on the synthetic Venice-shaped problem (1778 cameras / 993 923 points / 5 001 946 observations); every
step starts from the same initial parameters (device-to-device restore inside the timed region).
  value             whole-job LM iterations/s with the problem resident in HBM
  e2e               the same through sfmba_ba_solve_host (HOST buffers: pinned H2D of every input, D2H of cameras and
                    points inside the timed region)
  roofline          the implicit-Schur matvec K3 (matvec kernel + k_pcg_step = one PCG iteration), algorithmic bytes over
                    CUDA-event time; `traffic` only when profiles/r2_ncu_traffic.json was captured from these sources
  parity_vs_golden  the timed answer against the committed full-size oracle golden (every rank; a mismatch exits 3)
  c3, ba_weak_damping   (N = 1) BASELINE.json config C3 and the slow-PCG regime, each with its own CPU / parity figures
  cpu_baseline      the NumPy/SciPy oracle port timed on this host on the full problem (one LM iteration)
With N > 1 (torchrun) the points are sharded by observation count across ranks (strong scaling), cameras replicated; the
partial sums cross NVLink through peer windows inside the PCG kernels.
`--impl reference` times the oracle port (the reference is pure Python and cannot travel; its literal code cannot run beyond
~3 k observations) on the host cores, on the full problem.
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
GOLDEN = os.path.join(ROOT, 'tests', 'golden')
COST_RTOL, PARAM_TOL = 1e-6, 1e-5            # BASELINE.json north_star: the parity bar
CPU_BUDGET_S = 150.0                         # wall-clock cap of the CPU legs (repetitions are cut, never the problem)


def _peaks():
    try:
        with open(os.path.join(ROOT, 'MEASURED_PEAKS.json')) as f:
            return float(json.load(f)['hbm_gbs']), 'measured'
    except Exception:
        return 6650.0, 'fallback'


def source_sha():
    """Identity of the kernels: sha256 over the CUDA sources (what an ncu traffic record must have been taken from)."""
    import hashlib
    h = hashlib.sha256()
    d = os.path.join(ROOT, 'structure-from-motion_b200', 'csrc')
    for f in sorted(os.listdir(d)):
        with open(os.path.join(d, f), 'rb') as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def ncu_traffic(kernel, shape, scale, world):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the newest ncu capture (profiles/ncu_traffic.py writes
    profiles/r2_ncu_traffic.json).  Only used when that capture was taken from the sources this run is built from."""
    try:
        with open(os.path.join(ROOT, 'profiles', 'r2_ncu_traffic.json')) as f:
            rec = json.load(f)
        if rec.get('source_sha') != source_sha() or rec.get('shape') != shape or float(rec.get('scale', 1.0)) != scale \
                or int(rec.get('n_gpus', 1)) != world:
            return None, 'profiles/r2_ncu_traffic.json is from other sources / another shape: not quoted'
        return rec['kernels'].get(kernel), 'profiles/r2_ncu_traffic.json (%s)' % rec.get('report')
    except Exception:
        return None, 'no ncu capture on record'


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


def cpu_baseline(steps=1, warmup=0, shape='venice', scale=1.0):
    """The reference's CPU path for this workload: oracle/sfm_oracle.py (NumPy + scipy.sparse port of
    ba_processor.py:297-406 with the exact dense reduced solve) on the host cores, at the FULL shape.  One timed step is ONE
    LM iteration; repetitions are cut so that the leg ends within CPU_BUDGET_S (the problem never is)."""
    sys.path.insert(0, os.path.join(ROOT, 'oracle'))
    import sfm_oracle as oracle
    from sfm_b200 import synthetic
    prob, _ = synthetic.make_shape(shape, scale=scale)
    b = prob.obs_uv.T.astype(np.float64)
    times = []
    t_start = time.perf_counter()
    for s in range(warmup + steps):
        t0 = time.perf_counter()
        oracle.ba_solve(prob.cams, prob.pts, prob.obs_cam, prob.obs_pt, b, LAMBDA, 1, sparse=True, return_costs=False)
        dt = time.perf_counter() - t0
        if s >= warmup or not times:
            times.append(dt)
        if time.perf_counter() - t_start + dt > CPU_BUDGET_S:      # the next repetition would not fit
            break
    per_iter = float(np.mean(times))
    return {
        'value': 1.0 / per_iter, 'unit': 'LM iterations/s', 'cores': os.cpu_count(), 'kind': 'port',
        'sample': '%d LM iteration(s) of oracle/sfm_oracle.py (NumPy + scipy.sparse, exact dense reduced solve) on the full '
                  '%s problem (%d cams / %d pts / %d obs), %.1f s each; no extrapolation'
                  % (len(times), shape, prob.num_cams, prob.num_pts, prob.num_obs, per_iter),
        'obs_per_s': prob.num_obs / per_iter, 'ms_per_step': per_iter * 1e3, 'steps_run': len(times),
    }


def workload_name(shape, V, N, m, pcg_tol):
    return ('%s-shaped BA: %d cams / %d pts / %d obs; reference mode: lambda=%g, %d LM iterations per step, PCG to %.0e'
            % (shape, V, N, m, LAMBDA, LM_ITERS, pcg_tol))


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    base = cpu_baseline(steps=max(1, args.steps), warmup=args.warmup, shape=args.shape, scale=args.scale)
    from sfm_b200 import synthetic
    V, N, m, _ = synthetic.SHAPES[args.shape]
    line = {
        'impl': 'reference', 'metric': 'ba_lm_iterations_per_sec', 'value': base['value'], 'unit': 'LM iterations/s',
        'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': base['ms_per_step'],
        'higher_is_better': True, 'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
        'config': {'workload': workload_name(args.shape, V, N, m, args.pcg_tol),
                   'note': 'CPU arm: each timed step is ONE LM iteration of the oracle port (exact reduced solve instead of '
                           'PCG) on the full workload; %d of the %d requested steps were run (wall-clock cap %.0f s)'
                           % (base['steps_run'], args.steps, CPU_BUDGET_S),
                   'sample_scale': args.scale, 'steps_run': base['steps_run']},
        'cpu_baseline': {k: base[k] for k in ('value', 'unit', 'cores', 'kind', 'sample')},
        'e2e': {'value': base['value'], 'unit': 'LM iterations/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'jacobian_schur_obs_per_sec': base['obs_per_s'],
    }
    emit(line)


def max_rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(1.0, np.abs(b)))) if a.size else 0.0


def golden_parity(name, cams, pts, final_cost, pt_range=None):
    """Distance of a finished solve from the committed full-size oracle golden (tests/golden/make_golden_bal.py).
    pt_range = (p0, p1): this rank holds the points [p0, p1) only."""
    z = np.load(os.path.join(GOLDEN, name + '.npz'))
    idx, ref = z['pts_idx'], z['pts_sample']
    if pt_range is not None:
        keep = (idx >= pt_range[0]) & (idx < pt_range[1])
        idx, ref = idx[keep] - pt_range[0], ref[keep]
    return {'cams_max_rel': max_rel(cams, z['cams']), 'pts_max_rel': max_rel(pts[idx], ref),
            'cost_rel': abs(final_cost - float(z['costs'][-1])) / float(z['costs'][-1])}


def extra_c3(eng):
    """BASELINE.json config C3 in the same run: 1 M-point nonlinear triangulation (a18) and 10 k-view nonlinear PnP (a19)
    through the host-buffer entry points (copies inside the time), with the oracle timed on a sample of the same batch."""
    sys.path.insert(0, os.path.join(ROOT, 'oracle'))
    import sfm_oracle as oracle
    from sfm_b200 import synthetic

    def best_of(fn, n=2):
        fn()
        best = 1e30
        for _ in range(n):
            t0 = time.perf_counter()
            fn()
            best = min(best, time.perf_counter() - t0)
        return best
    out = {}
    M, it_tri = 1000000, 100
    projs, rowptr, obs_view, uv, X0 = synthetic.tri_problem(M, np.random.default_rng(3))
    dt = best_of(lambda: eng.triangulate_refine(projs, rowptr, obs_view, uv, X0, 0.5, it_tri))
    n_cpu = 300
    t0 = time.perf_counter()
    for j in range(n_cpu):
        o = slice(rowptr[j], rowptr[j + 1])
        pairs = [np.array([[uv[q, 0]], [uv[q, 1]], [1.0]]) for q in range(o.start, o.stop)]
        oracle.nonlinear_triangulate(np.append(X0[j], 1.0).reshape(4, 1), projs[obs_view[o]], pairs, 0.5, it_tri)
    cpu = (time.perf_counter() - t0) / (n_cpu * it_tri)
    out['triangulation'] = {
        'workload': '%d points, %d observations, lambda 0.5, %d iterations (triangulation_processor.py:160-234)' % (M, int(rowptr[-1]), it_tri),
        'seconds': dt, 'value': M * it_tri / dt, 'unit': 'point-iterations/s',
        'cpu_baseline': {'value': 1.0 / cpu, 'unit': 'point-iterations/s', 'cores': 1, 'kind': 'port',
                         'sample': 'oracle.nonlinear_triangulate on the first %d points of the batch' % n_cpu},
        'bound': 'latency / FP64 issue: one thread per point, the whole damped Gauss-Newton loop in registers; the batch '
                 'moves 48 B per point once (16 MB + 64 MB of observations), far below HBM time'}
    P, it_pnp = 10000, 300
    rp, xy, X, rots, locs = synthetic.pnp_problem(P, np.random.default_rng(4))
    dt = best_of(lambda: eng.pnp_refine(rp, xy, X, rots, locs, 5.0, it_pnp))
    n_cpu = 8
    t0 = time.perf_counter()
    for v in range(n_cpu):
        o = slice(rp[v], rp[v + 1])
        n = o.stop - o.start
        oracle.nonlinear_pnp(np.vstack((xy[o].T, np.ones(n))), np.vstack((X[o].T, np.ones(n))), np.eye(3), rots[v],
                             locs[v].reshape(3, 1), 5.0, it_pnp)
    cpu = (time.perf_counter() - t0) / (int(rp[n_cpu]) * it_pnp)
    out['pnp'] = {
        'workload': '%d views, %d observations, lambda 5, %d iterations (campose_processor.py:308-459)' % (P, int(rp[-1]), it_pnp),
        'seconds': dt, 'value': int(rp[-1]) * it_pnp / dt, 'unit': 'observation-iterations/s',
        'cpu_baseline': {'value': 1.0 / cpu, 'unit': 'observation-iterations/s', 'cores': 1, 'kind': 'port',
                         'sample': 'oracle.nonlinear_pnp on the first %d views of the batch' % n_cpu},
        'bound': 'latency / FP64 issue: one warp per view, observations re-read from L1/L2 every iteration'}
    return out


def extra_weak_damping(eng):
    """The matvec-dominated regime: lambda = 1e-2, exact dv/dC Jacobian, quarter-Venice, 2 LM iterations, default options
    (the configuration of tests/golden/bal_venice_weak.npz), device-resident."""
    import torch
    from sfm_b200 import synthetic
    z = np.load(os.path.join(GOLDEN, 'bal_venice_weak.npz'))
    prob, _ = synthetic.make_shape(str(z['shape']), scale=float(z['scale']))
    t = eng.to_device(prob)
    c0, p0 = t['cams'].clone(), t['pts'].clone()
    opt = eng.make_options(lam=float(z['lam']), iterations=int(z['iterations']), dvdc_sign_quirk=bool(z['dvdc_quirk']))
    best, res = 1e30, None
    for _ in range(5):
        t['cams'].copy_(c0)
        t['pts'].copy_(p0)
        torch.cuda.synchronize()
        res = eng.ba_solve_device(t, opt)
        best = min(best, res.ms['total'])
    par = golden_parity('bal_venice_weak', t['cams'].cpu().numpy(), t['pts'].cpu().numpy(), res.final_cost)
    its = int(z['iterations'])
    return {'workload': 'venice x %.2f (%d cams / %d pts / %d obs), lambda=%g, exact dv/dC, %d LM iterations, default options'
                        % (float(z['scale']), prob.num_cams, prob.num_pts, prob.num_obs, float(z['lam']), its),
            'value': its / (best * 1e-3), 'unit': 'LM iterations/s', 'device_ms': best,
            'pcg_iterations_per_lm_iteration': res.pcg_iterations / its,
            'pcg_fp64_refine_iterations': res.pcg_refine_iterations,
            'matvec_precision': {1: 'f64', 2: 'f32'}.get(res.matvec_precision), 'parity_vs_golden': par}


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
    pt_range = None
    if world > 1:
        sb = synthetic.shard_bounds(full.rowptr, world)
        pt_range = (int(sb[rank]), int(sb[rank + 1]))
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
    eng.profile_enable(args.profile_timed, phases=args.profile_timed)      # the timed region runs without any event spans
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
    ev1.record()
    barrier()
    ms = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device='cuda')
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    # ---- parity of the timed answer: every rank compares its cameras, its points and the global cost with the committed
    # full-size oracle golden of this exact configuration
    parity = None
    golden_name = {'venice': 'bal_venice_ref', 'ladybug': 'bal_ladybug_ref'}.get(args.shape)
    if golden_name and args.scale == 1.0 and args.pcg_max_iter == 0:
        par = golden_parity(golden_name, t['cams'].cpu().numpy(), t['pts'].cpu().numpy(), res.final_cost, pt_range)
        pv = torch.tensor([par['cams_max_rel'], par['pts_max_rel'], par['cost_rel']], dtype=torch.float64, device='cuda')
        if world > 1:
            dist.all_reduce(pv, op=dist.ReduceOp.MAX)
        parity = {'golden': 'tests/golden/%s.npz' % golden_name, 'cams_max_rel': float(pv[0]), 'pts_max_rel': float(pv[1]),
                  'cost_rel': float(pv[2]), 'ranks_compared': world,
                  'ok': bool(pv[0] <= PARAM_TOL and pv[1] <= PARAM_TOL and pv[2] <= COST_RTOL)}
    # ---- the same with sfmba_options.reuse_index (a caller that solves the same structure again keeps the observation
    # index): reported beside the headline, never as it
    opt_reuse = eng.make_options(lam=LAMBDA, iterations=LM_ITERS, pcg_tol=args.pcg_tol, compute_costs=True,
                                 pcg_max_iter=args.pcg_max_iter, reuse_index=True,
                                 matvec_precision={'auto': 0, 'f64': 1, 'f32': 2}[args.matvec])
    for _ in range(2):
        t['cams'].copy_(cams0)
        t['pts'].copy_(pts0)
        eng.ba_solve_device(t, opt_reuse)
    barrier()
    er0, er1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    er0.record()
    for _ in range(args.steps):
        t['cams'].copy_(cams0)
        t['pts'].copy_(pts0)
        res_reuse = eng.ba_solve_device(t, opt_reuse)
    er1.record()
    barrier()
    msr = torch.tensor([er0.elapsed_time(er1)], dtype=torch.float64, device='cuda')
    if world > 1:
        dist.all_reduce(msr, op=dist.ReduceOp.MAX)
    ms_reuse = float(msr.item())
    reuse_same = bool(res_reuse.final_cost == res.final_cost)
    # phase times: untimed steps with the phase spans on (no event records between the kernels of a phase)
    eng.profile_enable(False, phases=True)
    phase_steps = 3
    for _ in range(phase_steps):
        t['cams'].copy_(cams0)
        t['pts'].copy_(pts0)
        rp = eng.ba_solve_device(t, opt)
        for k in phase:
            phase[k] += rp.ms[k]
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
    prof = {k: eng.profile_get(k) for k in ('linearize', 'linearize32', 'matvec', 'pcg_tail', 'backsub', 'cost')}
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
    up_keys = ['cams', 'pts', 'obs_uv', 'obs_cam', 'rowptr']                 # obs_pt is not uploaded
    if V <= 65536 and not args.cam_i32:
        # half-width camera indices on the bus (SFMBA_PROBLEM_CAM_U16): prepared once, like the other pinned inputs
        pin['obs_cam16'] = pinned_like(prob.obs_cam.astype(np.uint16))
        hp.obs_cam16 = pin['obs_cam16'].numpy()
        up_keys[3] = 'obs_cam16'
    cams_h0, pts_h0 = prob.cams.copy(), prob.pts.copy()
    h2d = sum(getattr(hp, k).nbytes for k in up_keys)
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
        mvk = 'k_matvec32' if res.matvec_precision == 2 else 'k_matvec'
        k3_name = 'K3 implicit-Schur matvec = %s + k_pcg_step (one PCG iteration, rank 0 shard)' % mvk
        k3_traffic, k3_src = ncu_traffic(mvk, args.shape, args.scale, world)
        t64, k12_src = ncu_traffic('k_linearize', args.shape, args.scale, world)
        t32, _ = ncu_traffic('k_linearize32', args.shape, args.scale, world)
        n64, n32 = prof['linearize'][1], prof['linearize32'][1]
        k12_traffic = None                           # mean DRAM bytes per launch over the step's FP64 and FP32-Jacobian launches
        if (n64 == 0 or t64 is not None) and (n32 == 0 or t32 is not None) and n64 + n32 > 0:
            k12_traffic = ((t64 or 0.0) * n64 + (t32 or 0.0) * n32) / (n64 + n32)
        k3_gbs = k3_bytes / (k3_ms * 1e-3) / 1e9 if k3_ms > 0 else 0.0
        lin_ms = prof['linearize'][0] + prof['linearize32'][0]          # k_linearize (FP64) and k_linearize32 launches
        lin_n = prof['linearize'][1] + prof['linearize32'][1]
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
                       'matvec_precision': {1: 'f64', 2: 'f32'}.get(res.matvec_precision),
                       'matvec_precision_option': args.matvec,
                       'l2': 'inputs larger than L2 (>1 GB streamed per matvec)'},
            'jacobian_schur_obs_per_sec': (m / (k12_ms * 1e-3)) if k12_ms > 0 else None,
            'final_cost': res.final_cost, 'initial_cost': res.initial_cost,
            'phase_ms_per_step': {k: v / phase_steps for k, v in phase.items()},
            'kernel_ms': {k: (v[0] / max(prof_pcg_iters if k in ('matvec', 'pcg_tail') else v[1], 1))
                          for k, v in prof.items() if v[1] > 0},
            'roofline': {'kernel': k3_name,
                         'bound': 'hbm', 'achieved': k3_gbs, 'peak': peak, 'unit': 'GB/s', 'frac': k3_gbs / peak,
                         'peak_kind': peak_kind, 'traffic': k3_traffic, 'traffic_source': k3_src,
                         'algorithmic_bytes': k3_bytes, 'ms': k3_ms,
                         'matvec_alone': {'ms': mv_a_ms / max(prof_pcg_iters, 1),
                                          'frac': (k3_bytes / (mv_a_ms / max(prof_pcg_iters, 1) * 1e-3) / 1e9 / peak)
                                          if mv_a_ms > 0 else None}},
            'roofline_jacobian_schur': {'kernel': 'K1+K2 = k_linearize / k_linearize32 (residuals, Jacobians, point Schur blocks and camera Gram rows in one launch; J never stored); mean over the %d FP64 and %d FP32-Jacobian launches of a step' % (prof['linearize'][1], prof['linearize32'][1]), 'bound': 'hbm',
                                        'achieved': k12_gbs, 'peak': peak, 'unit': 'GB/s', 'frac': k12_gbs / peak,
                                        'algorithmic_bytes': k12_bytes, 'ms': k12_ms,
                                        'traffic': k12_traffic, 'traffic_source': k12_src},
            'e2e': {'value': LM_ITERS * e2e_steps / e2e_s, 'unit': 'LM iterations/s', 'h2d_bytes_per_step': h2d,
                    'd2h_bytes_per_step': d2h, 'steps': e2e_steps, 'ms_per_step': e2e_s / e2e_steps * 1e3,
                    'kernel_launches_per_step': launches_e2e, 'obs_cam_wire_format': 'u16' if 'obs_cam16' in up_keys else 'i32',
                    'pcie_gbs_if_serial': (h2d + d2h) / (e2e_s / e2e_steps) / 1e9},
            'gpu_launches': launches,
            'index_reused': {'value': LM_ITERS * args.steps / (ms_reuse * 1e-3), 'unit': 'LM iterations/s',
                             'ms_per_step': ms_reuse / args.steps, 'same_final_cost_bits': reuse_same,
                             'note': 'sfmba_options.reuse_index = 1: the caller keeps the observation index of the previous '
                                     'solve of the same structure; not the headline'},
            'clocks': sampler.summary(),
            'parity_vs_golden': parity,
        }
        if world == 1 and not args.no_extras:
            line['c3'] = extra_c3(eng)
            line['ba_weak_damping'] = extra_weak_damping(eng)
        if world == 1 and not args.no_cpu:
            base = cpu_baseline(steps=1, warmup=0, shape=args.shape, scale=args.scale)
            line['cpu_baseline'] = {k: base[k] for k in ('value', 'unit', 'cores', 'kind', 'sample')}
        emit(line)
    eng.close()
    if world > 1:
        dist.destroy_process_group()
    if parity is not None and not parity['ok']:
        sys.stderr.write('bench.py: the timed answer does not match the oracle golden: %r\n' % (parity,))
        sys.exit(3)


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
    ap.add_argument('--cam-i32', action='store_true', help='e2e: upload obs_cam as int32 instead of uint16')
    ap.add_argument('--no-extras', action='store_true', help='skip the c3 / ba_weak_damping keys (N = 1 only)')
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
