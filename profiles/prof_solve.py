#!/usr/bin/env python3
"""One BA solve at a named synthetic shape, for ncu (profiles/README: see DESIGN.md 'Evidence').
    python profiles/prof_solve.py [shape] [lm_iterations] [warm_solves] [auto|f64|f32]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sfm_b200                      # noqa: E402
from sfm_b200 import synthetic       # noqa: E402

shape = sys.argv[1] if len(sys.argv) > 1 else 'venice'
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 1
warm = int(sys.argv[3]) if len(sys.argv) > 3 else 0
prec = {'auto': 0, 'f64': 1, 'f32': 2}[sys.argv[4] if len(sys.argv) > 4 else 'auto']
eng = sfm_b200.Engine(device=0)
prob, _ = synthetic.make_shape(shape)
t = eng.to_device(prob)
c0, p0 = t['cams'].clone(), t['pts'].clone()
for _ in range(warm + 1):
    t['cams'].copy_(c0)
    t['pts'].copy_(p0)
    res = eng.ba_solve_device(t, lam=5.0, iterations=iters, matvec_precision=prec)
print('cost %.9g -> %.9g, %d PCG iterations, %d launches' % (res.initial_cost, res.final_cost, res.pcg_iterations,
                                                             res.kernel_launches))
eng.close()
