#!/usr/bin/env python3
"""CPU experiment behind `sfmba_options.precond`: PCG on the oracle's explicit reduced system with
the two block-Jacobi preconditioners -- the 7x7 blocks of A + lambda I against the diagonal blocks of the Schur complement
S -- counting the iterations to three tolerances.  Lives under tests/ because it drives oracle/ (test infrastructure); host only, not collected by pytest.
    python tests/precond_cpu_experiment.py > profiles/r1_precond_cpu_experiment.txt"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'oracle'))
import sfm_oracle as o               # noqa: E402
from sfm_b200 import synthetic       # noqa: E402

for shape, scale in (('venice', 0.1), ('ladybug', 1.0)):
    prob, _ = synthetic.make_shape(shape, scale=scale) if scale != 1.0 else synthetic.make_shape(shape)
    b = prob.obs_uv.T.astype(np.float64)
    V, N = prob.num_cams, prob.num_pts
    for lam in (5.0, 1e-2, 1e-4):
        ne = o.ba_normal_equations(prob.cams, prob.pts, prob.obs_cam, prob.obs_pt, b, lam)
        S, g, _ = o._reduced_system_sparse(ne, prob.obs_cam, prob.obs_pt, V, N)
        S = S.tocsr()
        Sb = S.tobsr(blocksize=(7, 7))
        Sd = np.zeros((V, 7, 7))
        for i in range(V):
            for k in range(Sb.indptr[i], Sb.indptr[i + 1]):
                if Sb.indices[k] == i:
                    Sd[i] = Sb.data[k]
        for name, Md in (('schur-jacobi', np.linalg.inv(Sd)), ('jacobi(A)', np.linalg.inv(ne['A']))):
            def prec(r):
                return np.einsum('vij,vj->vi', Md, r.reshape(V, 7)).reshape(-1)
            x = np.zeros(7 * V)
            r = g.copy()
            z = prec(r)
            p = z.copy()
            rz = r @ z
            gg = g @ g
            its = {}
            for it in range(1, 61):
                q = S @ p
                a = rz / (p @ q)
                x += a * p
                r -= a * q
                rr = r @ r
                for tol in (1e-6, 1e-8, 1e-10):
                    if rr <= tol * tol * gg and tol not in its:
                        its[tol] = it
                if len(its) == 3:
                    break
                z = prec(r)
                rzn = r @ z
                p = z + (rzn / rz) * p
                rz = rzn
            print('%s x %.2g  lambda %g  %-13s iterations to 1e-6 / 1e-8 / 1e-10: %s' %
                  (shape, scale, lam, name, ' / '.join(str(its.get(t, '>60')) for t in (1e-6, 1e-8, 1e-10))))
