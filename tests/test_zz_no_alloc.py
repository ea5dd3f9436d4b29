"""sfmba_ba_cost_noalloc (include/sfmba.h): the allocation-free variant of sfmba_ba_cost, whose scratch is a fixed
file-scope device array.  Collected last on purpose.

Checked through the C ABI against the existing entry point on the same device buffers, on two calls in a row with
different inputs (the scratch keeps the previous call's contents), plus the size limit and the bad-index flag."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-12        # same kernels, same summation order: the two entry points agree to the last bit or two


def _device_problem(torch, prob, dev):
    from sfm_b200 import _capi
    t = dict(obs_uv=torch.from_numpy(prob.obs_uv).to(dev), obs_cam=torch.from_numpy(prob.obs_cam).to(dev),
             obs_pt=torch.from_numpy(prob.obs_pt).to(dev), rowptr=torch.from_numpy(prob.rowptr).to(dev),
             cams=torch.from_numpy(prob.cams).to(dev), pts=torch.from_numpy(prob.pts).to(dev))
    cp = _capi.Problem()
    cp.num_cams, cp.num_pts, cp.num_obs = prob.num_cams, prob.num_pts, prob.num_obs
    cp.obs_uv, cp.obs_cam, cp.obs_pt = t['obs_uv'].data_ptr(), t['obs_cam'].data_ptr(), t['obs_pt'].data_ptr()
    cp.pt_rowptr, cp.cam_kinv = t['rowptr'].data_ptr(), None
    cp.cams, cp.pts = t['cams'].data_ptr(), t['pts'].data_ptr()
    return t, cp


def test_noalloc_cost_matches_ba_cost_twice(engine):
    import torch
    from sfm_b200 import _capi, synthetic
    lib = engine.lib
    dev = torch.device('cuda', engine.device)
    st = C.c_void_p(torch.cuda.current_stream(engine.device).cuda_stream)
    out = torch.full((1,), -1.0, dtype=torch.float64, device=dev)       # outputs are the caller's
    flags = torch.full((1,), 0xffff, dtype=torch.int32, device=dev)
    # two problems of different size back to back: the second call finds the first one's table and partials
    for shape, scale in (('ladybug', 1.0), ('small', 1.0)):
        prob, _ = synthetic.make_shape(shape, scale=scale)
        t, cp = _device_problem(torch, prob, dev)
        want = C.c_double(0.0)
        _capi.check(lib.sfmba_ba_cost(engine.ctx, C.byref(cp), C.byref(want), st), engine.ctx)
        rc = lib.sfmba_ba_cost_noalloc(engine.device, C.byref(cp), out.data_ptr(), flags.data_ptr(), st)
        assert rc == _capi.OK, _capi.last_error()
        torch.cuda.synchronize(engine.device)
        got = float(out.item())
        assert want.value > 0.0
        assert abs(got - want.value) <= TOL * want.value, (shape, got, want.value)
        assert int(flags.item()) == 0
        # and again on the same inputs after moving the points: still the existing entry point's answer
        t['pts'].add_(0.01)
        _capi.check(lib.sfmba_ba_cost(engine.ctx, C.byref(cp), C.byref(want), st), engine.ctx)
        rc = lib.sfmba_ba_cost_noalloc(engine.device, C.byref(cp), out.data_ptr(), None, st)
        assert rc == _capi.OK, _capi.last_error()
        torch.cuda.synchronize(engine.device)
        assert abs(float(out.item()) - want.value) <= TOL * want.value


def test_noalloc_cost_rejects_sizes_beyond_the_scratch(engine):
    import torch
    from sfm_b200 import _capi, synthetic
    lib = engine.lib
    dev = torch.device('cuda', engine.device)
    prob, _ = synthetic.make_shape('small')
    t, cp = _device_problem(torch, prob, dev)
    out = torch.zeros(1, dtype=torch.float64, device=dev)
    cp.num_cams = _capi.NOALLOC_MAX_CAMS + 1
    rc = lib.sfmba_ba_cost_noalloc(engine.device, C.byref(cp), out.data_ptr(), None, None)
    assert rc == _capi.ERR_UNSUPPORTED
    assert 'SFMBA_NOALLOC_MAX_CAMS' in _capi.last_error()
    cp.num_cams = prob.num_cams
    assert lib.sfmba_ba_cost_noalloc(engine.device, C.byref(cp), None, None, None) == _capi.ERR_INVALID_ARG


def test_noalloc_cost_flags_bad_index_and_clears_it_next_call(engine):
    import torch
    from sfm_b200 import _capi, synthetic
    lib = engine.lib
    dev = torch.device('cuda', engine.device)
    prob, _ = synthetic.make_shape('small')
    t, cp = _device_problem(torch, prob, dev)
    out = torch.zeros(1, dtype=torch.float64, device=dev)
    flags = torch.zeros(1, dtype=torch.int32, device=dev)
    good = int(t['obs_cam'][0].item())
    t['obs_cam'][0] = prob.num_cams + 5
    assert lib.sfmba_ba_cost_noalloc(engine.device, C.byref(cp), out.data_ptr(), flags.data_ptr(), None) == _capi.OK
    torch.cuda.synchronize(engine.device)
    assert int(flags.item()) & 64                           # SFMBA_FLAG_BAD_INDEX
    t['obs_cam'][0] = good
    assert lib.sfmba_ba_cost_noalloc(engine.device, C.byref(cp), out.data_ptr(), flags.data_ptr(), None) == _capi.OK
    torch.cuda.synchronize(engine.device)
    assert int(flags.item()) == 0                           # the shared flag word was cleared
    assert np.isfinite(float(out.item())) and float(out.item()) > 0.0
