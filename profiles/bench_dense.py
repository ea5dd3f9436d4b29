#!/usr/bin/env python3
"""Dense-Cholesky solver next to PCG, one JSON line per shape (device-resident, reference mode).

    python profiles/bench_dense.py [--shapes ladybug,venice:0.25,venice] [--iterations 3]

Per shape: device ms of a 3-iteration solve with each solver, the dense solver's assembly and
factor+solve kernel times (CUDA events inside the library) and the FP64 rate of the factorisation
((7V)^3 / 3 flops), and the largest difference between the two solvers' results.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sfm_b200                      # noqa: E402
from sfm_b200 import _capi, synthetic  # noqa: E402


def main():
    import torch
    ap = argparse.ArgumentParser()
    ap.add_argument('--shapes', default='ladybug,venice:0.25,venice')
    ap.add_argument('--iterations', type=int, default=3)
    ap.add_argument('--reps', type=int, default=3)
    args = ap.parse_args()
    eng = sfm_b200.Engine(device=0)
    for spec in args.shapes.split(','):
        name, _, scale = spec.partition(':')
        prob, _ = synthetic.make_shape(name, scale=float(scale) if scale else 1.0)
        t = eng.to_device(prob)
        c0, p0 = t['cams'].clone(), t['pts'].clone()
        out = {}
        for solver, label in ((_capi.SOLVER_PCG, 'pcg'), (_capi.SOLVER_DENSE_CHOLESKY, 'dense')):
            best, res = 1e30, None
            for rep in range(args.reps + 1):
                t['cams'].copy_(c0)
                t['pts'].copy_(p0)
                torch.cuda.synchronize()
                eng.profile_enable(rep == args.reps)
                eng.profile_reset()
                res = eng.ba_solve_device(t, lam=5.0, iterations=args.iterations, solver=solver, pcg_tol=1e-13)
                if rep < args.reps and rep > 0:
                    best = min(best, res.ms['total'])
            out[label] = dict(device_ms=best, final_cost=res.final_cost, phase_ms=res.ms,
                              cams=t['cams'].cpu().numpy().copy(), pts=t['pts'].cpu().numpy().copy())
            if label == 'dense':
                asm_ms, na = eng.profile_get('dense_assemble')
                fac_ms, nf = eng.profile_get('dense_factor_solve')
                n = 7 * prob.num_cams
                out[label]['assemble_ms'] = asm_ms / max(na, 1)
                out[label]['factor_solve_ms'] = fac_ms / max(nf, 1)
                out[label]['factor_tflops'] = (n ** 3 / 3.0) / (fac_ms / max(nf, 1) * 1e-3) / 1e12
            eng.profile_enable(False)
        dc = float(np.max(np.abs(out['pcg']['cams'] - out['dense']['cams'])))
        dp = float(np.max(np.abs(out['pcg']['pts'] - out['dense']['pts'])))
        for v in out.values():
            del v['cams'], v['pts']
        print(json.dumps({'workload': 'BA %s shape, %d LM iterations, lambda 5' % (spec, args.iterations),
                          'cams': prob.num_cams, 'pts': prob.num_pts, 'obs': prob.num_obs, 'reduced_dim': 7 * prob.num_cams,
                          'pcg_tol_1e-13': out['pcg'], 'dense_cholesky': out['dense'],
                          'max_abs_diff_cams': dc, 'max_abs_diff_pts': dp}))
        del t, c0, p0
        torch.cuda.empty_cache()
    eng.close()


if __name__ == '__main__':
    main()
