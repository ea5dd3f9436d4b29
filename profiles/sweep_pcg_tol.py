#!/usr/bin/env python3
"""PCG tolerance against iterations and against the distance from a tightly solved run (3 reference-mode LM
iterations at a named synthetic shape): the evidence behind the default `sfmba_options.pcg_tol`.
    python profiles/sweep_pcg_tol.py [shape]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sfm_b200                      # noqa: E402
from sfm_b200 import synthetic       # noqa: E402

shape = sys.argv[1] if len(sys.argv) > 1 else 'venice'
eng = sfm_b200.Engine(device=0)
prob, _ = synthetic.make_shape(shape)
t = eng.to_device(prob)
c0, p0 = t['cams'].clone(), t['pts'].clone()
ref = None
for tol in (1e-13, 1e-11, 1e-10, 1e-9, 1e-8, 1e-7, 1e-6, 1e-5):
    t['cams'].copy_(c0)
    t['pts'].copy_(p0)
    res = eng.ba_solve_device(t, lam=5.0, iterations=3, pcg_tol=tol)
    cams, pts = t['cams'].cpu().numpy().copy(), t['pts'].cpu().numpy().copy()
    if ref is None:
        ref = (cams, pts, res.final_cost)
    print(json.dumps({'shape': shape, 'pcg_tol': tol, 'pcg_iterations': res.pcg_iterations,
                      'final_cost': res.final_cost,
                      'cost_rel_diff': abs(res.final_cost - ref[2]) / ref[2],
                      'cams_max_diff': float(np.max(np.abs(cams - ref[0]) / np.maximum(1.0, np.abs(ref[0])))),
                      'pts_max_diff': float(np.max(np.abs(pts - ref[1]) / np.maximum(1.0, np.abs(ref[1]))))}))
eng.close()
