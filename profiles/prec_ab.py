#!/usr/bin/env python3
"""Matvec precision A/B (sfmba_options.matvec_precision): distance of the F64 / F32 / AUTO solves from the full-size oracle
goldens (tests/golden/bal_*.npz) and their device time.  One JSON line per (case, mode).

    python profiles/prec_ab.py [bal_venice_ref bal_venice_weak bal_ladybug_ref]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sfm_b200                      # noqa: E402
from sfm_b200 import _capi, synthetic  # noqa: E402


def max_rel(a, b):
    return float(np.max(np.abs(a - b) / np.maximum(1.0, np.abs(b))))


def main():
    import torch
    names = sys.argv[1:] or ['bal_venice_ref', 'bal_venice_weak', 'bal_ladybug_ref']
    eng = sfm_b200.Engine(device=0)
    eng.profile_enable(False, phases=True)
    for name in names:
        z = np.load(os.path.join(ROOT, 'tests', 'golden', name + '.npz'))
        prob, _ = synthetic.make_shape(str(z['shape']), scale=float(z['scale']))
        t = eng.to_device(prob)
        c0, p0 = t['cams'].clone(), t['pts'].clone()
        for mode, mname in ((_capi.MATVEC_F64, 'f64'), (_capi.MATVEC_F32, 'f32'), (_capi.MATVEC_AUTO, 'auto')):
            opt = eng.make_options(lam=float(z['lam']), iterations=int(z['iterations']),
                                   dvdc_sign_quirk=bool(z['dvdc_quirk']), matvec_precision=mode)
            best = 1e30
            for _ in range(4):
                t['cams'].copy_(c0)
                t['pts'].copy_(p0)
                torch.cuda.synchronize()
                res = eng.ba_solve_device(t, opt)
                best = min(best, res.ms['total'])
            cams, pts = t['cams'].cpu().numpy(), t['pts'].cpu().numpy()
            print(json.dumps({
                'case': name, 'mode': mname, 'ran': {1: 'f64', 2: 'f32'}.get(res.matvec_precision), 'device_ms': round(best, 4),
                'pcg_ms': round(res.ms['pcg'], 4), 'pcg_iterations': res.pcg_iterations,
                'refine_iterations': res.pcg_refine_iterations, 'relinearized': res.relinearized, 'flags': res.flags,
                'cams_vs_oracle': max_rel(cams, z['cams']), 'pts_vs_oracle': max_rel(pts[z['pts_idx']], z['pts_sample']),
                'cost_vs_oracle': abs(res.final_cost - float(z['costs'][-1])) / float(z['costs'][-1])}), flush=True)
    eng.close()


if __name__ == '__main__':
    main()
