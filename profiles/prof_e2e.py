#!/usr/bin/env python3
"""Host-buffer (end-to-end) solve time at the Venice shape for several upload-range counts, next to a plain
pinned H2D / D2H of the same bytes.   SFMBA_UPLOAD_RANGES is read per call.
    python profiles/prof_e2e.py"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sfm_b200                      # noqa: E402
from sfm_b200 import synthetic       # noqa: E402

eng = sfm_b200.Engine(device=0)
prob, _ = synthetic.make_shape('venice')


def pinned_like(a):
    t = torch.empty(a.shape, dtype=torch.from_numpy(a).dtype, pin_memory=True)
    t.numpy()[...] = a
    return t


pin = {k: pinned_like(getattr(prob, k)) for k in ('cams', 'pts', 'obs_uv', 'obs_cam', 'obs_pt', 'rowptr')}
hp = sfm_b200.BAProblem(cams=pin['cams'].numpy(), pts=pin['pts'].numpy(), obs_uv=pin['obs_uv'].numpy(),
                        obs_cam=pin['obs_cam'].numpy(), obs_pt=pin['obs_pt'].numpy(), rowptr=pin['rowptr'].numpy())
c0, p0 = prob.cams.copy(), prob.pts.copy()
# raw copies of the same bytes
dev = {k: torch.empty_like(v, device='cuda') for k, v in pin.items() if k != 'obs_pt'}
for _ in range(3):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for k in dev:
        dev[k].copy_(pin[k], non_blocking=True)
    torch.cuda.synchronize()
    t_h2d = time.perf_counter() - t0
    t0 = time.perf_counter()
    pin['pts'].copy_(dev['pts'], non_blocking=True)
    torch.cuda.synchronize()
    t_d2h = time.perf_counter() - t0
h2d_bytes = sum(v.numel() * v.element_size() for k, v in pin.items() if k != 'obs_pt')
print(json.dumps({'h2d_ms': t_h2d * 1e3, 'h2d_GBps': h2d_bytes / t_h2d / 1e9, 'd2h_ms_points': t_d2h * 1e3,
                  'd2h_GBps': pin['pts'].numel() * 8 / t_d2h / 1e9}))
for R in (1, 2, 4, 8, 4, 1):
    os.environ['SFMBA_UPLOAD_RANGES'] = str(R)
    ts = []
    for rep in range(7):
        hp.cams[...] = c0
        hp.pts[...] = p0
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        res = eng.ba_solve_host(hp, lam=5.0, iterations=3)
        ts.append(time.perf_counter() - t0)
    print(json.dumps({'upload_ranges': R, 'e2e_ms_min': min(ts[2:]) * 1e3, 'e2e_ms_median': float(np.median(ts[2:])) * 1e3,
                      'device_ms': res.ms['total'], 'final_cost': res.final_cost}))
