#!/usr/bin/env python3
"""Golden vectors of the BENCHMARKED bundle-adjustment configurations (BASELINE.json configs C4 / C2 and a
weak-damping case), from the CPU oracle at full size.

TEST INFRASTRUCTURE.  Run once in the build container (minutes of CPU, ~20 GB of RAM for full Venice):

    python tests/golden/make_golden_bal.py [bal_venice_ref] [bal_venice_weak] [bal_ladybug_ref]

The oracle (oracle/sfm_oracle.py, pinned to the unmodified reference by tests/test_oracle_golden.py) solves the
reduced camera system EXACTLY (dense LU of the 7V x 7V Schur complement, ba_processor.py:382) -- the GPU path solves
it by PCG to sfmba_options.pcg_tol, so these files are what justifies that tolerance on the headline shape.

Each <name>.npz holds: the generator arguments, checksums of the generated inputs (the test re-generates the problem
on the GPU box and refuses to compare if they differ), the cameras after the last iteration (V, 7), the cost history
(start of every iteration + final), a strided sample of the final points with its indices, and f64 checksums
(sum, sum of squares per coordinate) of ALL final points.
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'oracle'))

CASES = {
    # name: (shape, scale, lambda, iterations, dvdc_quirk)
    'bal_venice_ref': ('venice', 1.0, 5.0, 3, True),          # BASELINE.json C4 = bench.py's workload
    'bal_venice_weak': ('venice', 0.25, 1e-2, 2, False),      # weak damping, exact Jacobian: PCG needs 15-25 iterations
    'bal_ladybug_ref': ('ladybug', 1.0, 5.0, 3, True),        # BASELINE.json C2
}
PT_STRIDE = 997


def input_checksums(prob):
    return np.array([prob.cams.sum(), prob.pts.sum(), float(prob.obs_uv.astype(np.float64).sum()),
                     float(prob.obs_cam.astype(np.int64).sum()), float(prob.rowptr.astype(np.int64).sum())])


def main(names):
    import sfm_oracle as oracle
    from sfm_b200 import synthetic
    for name in names:
        shape, scale, lam, iters, quirk = CASES[name]
        prob, _ = synthetic.make_shape(shape, scale=scale)
        b = prob.obs_uv.T.astype(np.float64)
        t0 = time.time()
        cams, pts, costs = oracle.ba_solve(prob.cams, prob.pts, prob.obs_cam, prob.obs_pt, b, lam, iters,
                                           dvdc_quirk=quirk, sparse=True)
        dt = time.time() - t0
        idx = np.arange(0, prob.num_pts, PT_STRIDE)
        np.savez_compressed(
            os.path.join(HERE, name + '.npz'), shape=shape, scale=scale, lam=lam, iterations=iters, dvdc_quirk=quirk,
            sizes=np.array([prob.num_cams, prob.num_pts, prob.num_obs]), input_checksums=input_checksums(prob),
            cams=cams, costs=np.array(costs), pts_idx=idx, pts_sample=pts[idx],
            pts_sum=pts.sum(axis=0), pts_sumsq=(pts * pts).sum(axis=0), oracle_seconds=dt)
        print('%s: %d cams / %d pts / %d obs, %d iterations in %.1f s, costs %s' %
              (name, prob.num_cams, prob.num_pts, prob.num_obs, iters, dt, ['%.9g' % c for c in costs]), flush=True)


if __name__ == '__main__':
    main(sys.argv[1:] or list(CASES))
