"""Host-side engine: owns one sfmba_ctx per process/GPU and moves arrays across the C ABI.

PyTorch is used for device memory and streams only (tensor hand-off, `data_ptr()`); numpy for host
arrays.  Every compute call goes through libsfmba.so -- there is no eager / CPU fallback here.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from . import _capi


@dataclass
class BAProblem:
    """One bundle-adjustment problem in host memory, laid out as the C ABI wants it.

    cams (V,7) f64 [C, q(w,x,y,z)]; pts (N,3) f64; obs_uv (2,m) f32 SoA; obs_cam / obs_pt (m,) i32 in
    point-major, view-minor order (ba_processor.py:304-306); rowptr (N+1,) i32; kinv (V,9) f64 or None.
    """
    cams: np.ndarray
    pts: np.ndarray
    obs_uv: np.ndarray
    obs_cam: np.ndarray
    obs_pt: np.ndarray
    rowptr: np.ndarray
    kinv: np.ndarray | None = None
    obs_key: np.ndarray | None = None      # key-point index of each observation (bookkeeping only)
    obs_cam16: np.ndarray | None = None    # obs_cam as uint16 (V <= 65536): the host-buffer solve uploads this instead

    def __post_init__(self):
        self.cams = np.ascontiguousarray(self.cams, dtype=np.float64).reshape(-1, 7)
        self.pts = np.ascontiguousarray(self.pts, dtype=np.float64).reshape(-1, 3)
        self.obs_uv = np.ascontiguousarray(self.obs_uv, dtype=np.float32).reshape(2, -1)
        self.obs_cam = np.ascontiguousarray(self.obs_cam, dtype=np.int32).reshape(-1)
        self.obs_pt = np.ascontiguousarray(self.obs_pt, dtype=np.int32).reshape(-1)
        self.rowptr = np.ascontiguousarray(self.rowptr, dtype=np.int32).reshape(-1)
        if self.kinv is not None:
            self.kinv = np.ascontiguousarray(self.kinv, dtype=np.float64).reshape(-1, 9)
        m = self.obs_cam.shape[0]
        if self.obs_uv.shape[1] != m or self.obs_pt.shape[0] != m:
            raise ValueError('observation arrays disagree in length')
        if self.rowptr.shape[0] != self.pts.shape[0] + 1 or (m and int(self.rowptr[-1]) != m):
            raise ValueError('rowptr does not describe the observations')
        if self.kinv is not None and self.kinv.shape[0] != self.cams.shape[0]:
            raise ValueError('kinv needs one row per camera')
        if m and (self.obs_cam.min() < 0 or self.obs_cam.max() >= self.cams.shape[0]):
            raise ValueError('camera index out of range')

    @property
    def num_cams(self):
        return self.cams.shape[0]

    @property
    def num_pts(self):
        return self.pts.shape[0]

    @property
    def num_obs(self):
        return self.obs_cam.shape[0]


@dataclass
class SolveResult:
    initial_cost: float
    final_cost: float
    final_lambda: float
    iterations_done: int
    iterations_accepted: int
    pcg_iterations: int
    flags: int
    costs: list = field(default_factory=list)
    ms: dict = field(default_factory=dict)
    kernel_launches: int = 0
    pcg_refine_iterations: int = 0
    matvec_precision: int = 0
    relinearized: int = 0


def _ptr(a):
    if a is None:
        return None
    return a.ctypes.data_as(C.c_void_p)


def _raise_numeric(err: _capi.SfmbaError, flags: int):
    """Map device numeric flags onto the exceptions the reference raises."""
    if flags & _capi.FLAG_QW_NEAR_ZERO:
        # utils.py:49-51
        raise ValueError('convert_rotation_to_quaternion : Invalid output qw (|qw| < 1e-6)') from err
    raise FloatingPointError(str(err)) from err


class Engine:
    """One libsfmba context bound to one CUDA device."""

    def __init__(self, device: int = 0, world_size: int = 1, rank: int = 0, nccl_unique_id: bytes | None = None):
        self.lib = _capi.load_library()
        self.ctx = C.c_void_p()
        idbuf = None
        if world_size > 1:
            if nccl_unique_id is None or len(nccl_unique_id) != 128:
                raise ValueError('world_size > 1 needs the 128-byte ncclUniqueId of rank 0')
            idbuf = C.create_string_buffer(bytes(nccl_unique_id), 128)
        rc = self.lib.sfmba_create(C.byref(self.ctx), device, world_size, rank,
                                   C.cast(idbuf, C.c_void_p) if idbuf is not None else None)
        _capi.check(rc)
        self.device = device
        self.world_size = world_size
        self.rank = rank

    # -- lifetime ---------------------------------------------------------------------------------
    def close(self):
        if getattr(self, 'ctx', None) is not None and self.ctx:
            self.lib.sfmba_destroy(self.ctx)
            self.ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @staticmethod
    def nccl_unique_id() -> bytes:
        lib = _capi.load_library()
        buf = C.create_string_buffer(128)
        _capi.check(lib.sfmba_nccl_unique_id(C.cast(buf, C.c_void_p)))
        return buf.raw

    # -- options ----------------------------------------------------------------------------------
    def make_options(self, lam=5.0, iterations=3, mode=_capi.MODE_REFERENCE_FIXED, pcg_tol=1e-8, pcg_max_iter=0,
                     dvdc_sign_quirk=True, force_slow_path=False, compute_costs=True, solver=_capi.SOLVER_PCG,
                     huber_delta=0.0, precond=_capi.PRECOND_JACOBI, matvec_precision=_capi.MATVEC_AUTO, reuse_index=False):
        opt = _capi.default_options()
        opt.solver = int(solver)
        opt.lambda_ = float(lam)
        opt.iterations = int(iterations)
        opt.mode = int(mode)
        opt.pcg_tol = float(pcg_tol)
        opt.pcg_max_iter = int(pcg_max_iter)
        opt.dvdc_sign_quirk = 1 if dvdc_sign_quirk else 0
        opt.force_slow_path = 1 if force_slow_path else 0
        opt.compute_costs = 1 if compute_costs else 0
        opt.huber_delta = float(huber_delta)
        opt.precond = int(precond)
        opt.matvec_precision = int(matvec_precision)
        opt.reuse_index = 1 if reuse_index else 0
        return opt

    @staticmethod
    def _result(res: _capi.Result) -> SolveResult:
        return SolveResult(res.initial_cost, res.final_cost, res.final_lambda, res.iterations_done,
                           res.iterations_accepted, res.pcg_iterations, res.flags,
                           [res.costs[i] for i in range(res.num_costs)],
                           dict(total=res.ms_total, index=res.ms_index, linearize=res.ms_linearize,
                                camreduce=res.ms_camreduce, pcg=res.ms_pcg, update=res.ms_update),
                           res.kernel_launches, res.pcg_refine_iterations, res.matvec_precision, res.relinearized)

    # -- BA: host buffers (the end-to-end path) ---------------------------------------------------
    def _host_problem(self, p: BAProblem) -> _capi.Problem:
        cp = _capi.Problem()
        cp.num_cams, cp.num_pts, cp.num_obs = p.num_cams, p.num_pts, p.num_obs
        cp.obs_uv, cp.obs_cam, cp.obs_pt = _ptr(p.obs_uv), _ptr(p.obs_cam), _ptr(p.obs_pt)
        cp.pt_rowptr, cp.cam_kinv = _ptr(p.rowptr), _ptr(p.kinv)
        cp.cams, cp.pts = _ptr(p.cams), _ptr(p.pts)
        if p.obs_cam16 is not None:
            if p.obs_cam16.dtype != np.uint16 or p.obs_cam16.shape != p.obs_cam.shape or not p.obs_cam16.flags.c_contiguous:
                raise ValueError('obs_cam16 must be a contiguous uint16 copy of obs_cam')
            cp.obs_cam = _ptr(p.obs_cam16)
            cp.flags = _capi.PROBLEM_CAM_U16
        return cp

    def ba_solve_host(self, p: BAProblem, opt: _capi.Options | None = None, **kw) -> SolveResult:
        """Solve in place: p.cams and p.pts are overwritten with the refined parameters."""
        if opt is None:
            opt = self.make_options(**kw)
        cp = self._host_problem(p)
        res = _capi.Result()
        rc = self.lib.sfmba_ba_solve_host(self.ctx, C.byref(cp), C.byref(opt), C.byref(res))
        if rc == _capi.ERR_NUMERIC:
            try:
                _capi.check(rc, self.ctx)
            except _capi.SfmbaError as e:
                _raise_numeric(e, res.flags)
        _capi.check(rc, self.ctx)
        return self._result(res)

    # -- BA: device buffers (torch tensors) --------------------------------------------------------
    def to_device(self, p: BAProblem):
        """Copy a host problem into torch CUDA tensors; returns a dict usable with ba_solve_device."""
        import torch
        dev = torch.device('cuda', self.device)
        t = dict(
            obs_uv=torch.from_numpy(p.obs_uv).to(dev), obs_cam=torch.from_numpy(p.obs_cam).to(dev),
            obs_pt=torch.from_numpy(p.obs_pt).to(dev), rowptr=torch.from_numpy(p.rowptr).to(dev),
            kinv=torch.from_numpy(p.kinv).to(dev) if p.kinv is not None else None,
            cams=torch.from_numpy(p.cams).to(dev), pts=torch.from_numpy(p.pts).to(dev),
            V=p.num_cams, N=p.num_pts, m=p.num_obs)
        return t

    @staticmethod
    def _dev_problem(t) -> _capi.Problem:
        cp = _capi.Problem()
        cp.num_cams, cp.num_pts, cp.num_obs = t['V'], t['N'], t['m']
        cp.obs_uv, cp.obs_cam, cp.obs_pt = t['obs_uv'].data_ptr(), t['obs_cam'].data_ptr(), t['obs_pt'].data_ptr()
        cp.pt_rowptr = t['rowptr'].data_ptr()
        cp.cam_kinv = t['kinv'].data_ptr() if t.get('kinv') is not None else None
        cp.cams, cp.pts = t['cams'].data_ptr(), t['pts'].data_ptr()
        return cp

    def ba_solve_device(self, t, opt: _capi.Options | None = None, stream=None, **kw) -> SolveResult:
        import torch
        if opt is None:
            opt = self.make_options(**kw)
        if stream is None:
            stream = torch.cuda.current_stream(self.device).cuda_stream
        cp = self._dev_problem(t)
        res = _capi.Result()
        rc = self.lib.sfmba_ba_solve(self.ctx, C.byref(cp), C.byref(opt), C.byref(res), C.c_void_p(stream))
        if rc == _capi.ERR_NUMERIC:
            try:
                _capi.check(rc, self.ctx)
            except _capi.SfmbaError as e:
                _raise_numeric(e, res.flags)
        _capi.check(rc, self.ctx)
        return self._result(res)

    def residual_jacobian(self, p: BAProblem, dvdc_sign_quirk=True, want_r64=True):
        """K1 alone.  Returns (r (m,2) f32, Jp (m,2,7) f32, Jx (m,2,3) f32, r64 (m,2) f64 | None, cost)."""
        import torch
        t = self.to_device(p)
        dev = torch.device('cuda', self.device)
        m = p.num_obs
        r = torch.empty((2, max(m, 1)), dtype=torch.float32, device=dev)
        jp = torch.empty((14, max(m, 1)), dtype=torch.float32, device=dev)
        jx = torch.empty((6, max(m, 1)), dtype=torch.float32, device=dev)
        r64 = torch.empty((2, max(m, 1)), dtype=torch.float64, device=dev) if want_r64 else None
        cost = torch.zeros(1, dtype=torch.float64, device=dev)
        cp = self._dev_problem(t)
        st = torch.cuda.current_stream(self.device).cuda_stream
        rc = self.lib.sfmba_residual_jacobian(self.ctx, C.byref(cp), 1 if dvdc_sign_quirk else 0, r.data_ptr(),
                                              jp.data_ptr(), jx.data_ptr(), r64.data_ptr() if want_r64 else None,
                                              cost.data_ptr(), C.c_void_p(st))
        _capi.check(rc, self.ctx)
        torch.cuda.synchronize(self.device)
        r_h = r[:, :m].t().cpu().numpy()
        jp_h = jp[:, :m].t().reshape(m, 2, 7).cpu().numpy()
        jx_h = jx[:, :m].t().reshape(m, 2, 3).cpu().numpy()
        r64_h = r64[:, :m].t().cpu().numpy() if want_r64 else None
        return r_h, jp_h, jx_h, r64_h, float(cost.item())

    def residual_jacobian_f64(self, p: BAProblem, dvdc_sign_quirk=True):
        """K1 alone in full precision.  Returns (r (m,2), Jp (m,2,7), Jx (m,2,3), cost), all float64."""
        import torch
        t = self.to_device(p)
        dev = torch.device('cuda', self.device)
        m = p.num_obs
        r = torch.empty((2, max(m, 1)), dtype=torch.float64, device=dev)
        jp = torch.empty((14, max(m, 1)), dtype=torch.float64, device=dev)
        jx = torch.empty((6, max(m, 1)), dtype=torch.float64, device=dev)
        cost = torch.zeros(1, dtype=torch.float64, device=dev)
        cp = self._dev_problem(t)
        st = torch.cuda.current_stream(self.device).cuda_stream
        rc = self.lib.sfmba_residual_jacobian_f64(self.ctx, C.byref(cp), 1 if dvdc_sign_quirk else 0, r.data_ptr(),
                                                  jp.data_ptr(), jx.data_ptr(), cost.data_ptr(), C.c_void_p(st))
        _capi.check(rc, self.ctx)
        return (r[:, :m].t().cpu().numpy(), jp[:, :m].t().reshape(m, 2, 7).cpu().numpy(),
                jx[:, :m].t().reshape(m, 2, 3).cpu().numpy(), float(cost.item()))

    def ba_cost(self, p: BAProblem) -> float:
        import torch
        t = self.to_device(p)
        cp = self._dev_problem(t)
        out = C.c_double(0.0)
        st = torch.cuda.current_stream(self.device).cuda_stream
        _capi.check(self.lib.sfmba_ba_cost(self.ctx, C.byref(cp), C.byref(out), C.c_void_p(st)), self.ctx)
        return out.value

    # -- refiners (host buffers) -------------------------------------------------------------------
    def triangulate_refine(self, projs, rowptr, obs_view, obs_uv, pts, lam, iterations):
        """a18 batched.  projs (P,3,4); rowptr (M+1,); obs_view (m,); obs_uv (m,2); pts (M,3).
        Returns refined (M,3) f64."""
        projs = np.ascontiguousarray(projs, dtype=np.float64).reshape(-1, 12)
        rowptr = np.ascontiguousarray(rowptr, dtype=np.int32)
        obs_view = np.ascontiguousarray(obs_view, dtype=np.int32)
        obs_uv = np.ascontiguousarray(obs_uv, dtype=np.float64).reshape(-1, 2)
        out = np.array(pts, dtype=np.float64, order='C', copy=True).reshape(-1, 3)
        flags = C.c_uint32(0)
        rc = self.lib.sfmba_triangulate_refine_batched_host(
            self.ctx, out.shape[0], obs_view.shape[0], projs.shape[0], _ptr(projs), _ptr(rowptr), _ptr(obs_view),
            _ptr(obs_uv), _ptr(out), float(lam), int(iterations), C.byref(flags))
        if rc == _capi.ERR_NUMERIC:
            raise FloatingPointError('triangulation produced non-finite values (flags 0x%x)' % flags.value)
        _capi.check(rc, self.ctx)
        return out

    def triangulate_linear(self, projs, rowptr, obs_view, obs_uv):
        """f2 batched (triangulation_processor.py:136-155, campose_processor.py:181-189).  projs (P,3,4); rowptr
        (M+1,); obs_view (m,); obs_uv (m,2).  Returns (pts (M,3) f64, front (M,) bool: positive depth in
        every view of the point)."""
        projs = np.ascontiguousarray(projs, dtype=np.float64).reshape(-1, 12)
        rowptr = np.ascontiguousarray(rowptr, dtype=np.int32)
        obs_view = np.ascontiguousarray(obs_view, dtype=np.int32)
        obs_uv = np.ascontiguousarray(obs_uv, dtype=np.float64).reshape(-1, 2)
        M = rowptr.shape[0] - 1
        out = np.zeros((M, 3), dtype=np.float64)
        front = np.zeros(M, dtype=np.int32)
        flags = C.c_uint32(0)
        rc = self.lib.sfmba_triangulate_linear_batched_host(
            self.ctx, M, obs_view.shape[0], projs.shape[0], _ptr(projs), _ptr(rowptr), _ptr(obs_view), _ptr(obs_uv),
            _ptr(out), _ptr(front), C.byref(flags))
        if rc == _capi.ERR_NUMERIC:
            raise FloatingPointError('linear triangulation produced a point at infinity (flags 0x%x)' % flags.value)
        _capi.check(rc, self.ctx)
        return out, front.astype(bool)

    def pnp_refine(self, rowptr, obs_xy, obs_X, rot, loc, lam, iterations, row_packing_quirk=True,
                   dvdc_sign_quirk=True):
        """a19 batched.  rowptr (P+1,); obs_xy (m,2) normalised; obs_X (m,3); rot (P,3,3); loc (P,3).
        Returns (rot (P,3,3), loc (P,3))."""
        rowptr = np.ascontiguousarray(rowptr, dtype=np.int32)
        obs_xy = np.ascontiguousarray(obs_xy, dtype=np.float64).reshape(-1, 2)
        obs_X = np.ascontiguousarray(obs_X, dtype=np.float64).reshape(-1, 3)
        rot_o = np.array(rot, dtype=np.float64, order='C', copy=True).reshape(-1, 9)
        loc_o = np.array(loc, dtype=np.float64, order='C', copy=True).reshape(-1, 3)
        flags = C.c_uint32(0)
        rc = self.lib.sfmba_pnp_refine_batched_host(
            self.ctx, rot_o.shape[0], obs_xy.shape[0], _ptr(rowptr), _ptr(obs_xy), _ptr(obs_X), _ptr(rot_o),
            _ptr(loc_o), float(lam), int(iterations), 1 if row_packing_quirk else 0, 1 if dvdc_sign_quirk else 0,
            C.byref(flags))
        if rc == _capi.ERR_NUMERIC:
            if flags.value & _capi.FLAG_QW_NEAR_ZERO:
                raise ValueError('convert_rotation_to_quaternion : Invalid output qw (|qw| < 1e-6)')
            raise FloatingPointError('PnP refinement produced non-finite values (flags 0x%x)' % flags.value)
        _capi.check(rc, self.ctx)
        return rot_o.reshape(-1, 3, 3), loc_o

    # -- visibility index on the device ------------------------------------------------------------
    def pnp_ransac(self, key, X, k_mat, samples, inlier_threshold, loc_sign=None, want_all=False):
        """f3: every 6-point DLT hypothesis of campose_processor.py:485-633 on the device.  key (n,3), X (n,4),
        samples (H,6) int, loc_sign (H,) of +-1 or None.  Returns dict(hyp, count, mask (n,) bool, rot (3,3),
        loc (3,1), flags[, counts (H,), models (H,12)])."""
        key = np.ascontiguousarray(key, dtype=np.float64).reshape(-1, 3)
        X = np.ascontiguousarray(X, dtype=np.float64).reshape(-1, 4)
        k_mat = np.ascontiguousarray(k_mat, dtype=np.float64).reshape(9)
        samples = np.ascontiguousarray(samples, dtype=np.int32).reshape(-1, 6)
        n, H = key.shape[0], samples.shape[0]
        if X.shape[0] != n:
            raise ValueError('key / X size mismatch')
        if H and (samples.min() < 0 or samples.max() >= n):
            raise ValueError('sample index out of range')
        sign = None if loc_sign is None else np.ascontiguousarray(loc_sign, dtype=np.int32).reshape(H)
        rot = np.zeros(9)
        loc = np.zeros(3)
        best = np.zeros(2, dtype=np.int32)
        mask = np.zeros(max(n, 1), dtype=np.uint8)
        counts = np.zeros(max(H, 1), dtype=np.int32) if want_all else None
        models = np.zeros((max(H, 1), 12)) if want_all else None
        flags = C.c_uint32(0)
        rc = self.lib.sfmba_pnp_ransac_host(
            self.ctx, n, H, _ptr(key), _ptr(X), _ptr(k_mat), _ptr(samples), None if sign is None else _ptr(sign),
            float(inlier_threshold), _ptr(rot), _ptr(loc), _ptr(best), _ptr(mask),
            None if counts is None else _ptr(counts), None if models is None else _ptr(models), C.byref(flags))
        _capi.check(rc, self.ctx)
        out = dict(hyp=int(best[0]), count=int(best[1]), mask=mask[:n].astype(bool), rot=rot.reshape(3, 3),
                   loc=loc.reshape(3, 1), flags=int(flags.value))
        if want_all:
            out.update(counts=counts[:H], models=models[:H])
        return out

    def fund_ransac(self, pairs, samples, inlier_threshold, want_all=False):
        """f3: every 8-point hypothesis of epipolar_processor.py:140-247 on the device.  pairs (n,4) normalised
        [xl, yl, xr, yr], samples (H,8).  Returns dict(hyp, count, mask, F (3,3), flags[, counts, models (H,9)])."""
        pairs = np.ascontiguousarray(pairs, dtype=np.float64).reshape(-1, 4)
        samples = np.ascontiguousarray(samples, dtype=np.int32).reshape(-1, 8)
        n, H = pairs.shape[0], samples.shape[0]
        if H and (samples.min() < 0 or samples.max() >= n):
            raise ValueError('sample index out of range')
        F = np.zeros(9)
        best = np.zeros(2, dtype=np.int32)
        mask = np.zeros(max(n, 1), dtype=np.uint8)
        counts = np.zeros(max(H, 1), dtype=np.int32) if want_all else None
        models = np.zeros((max(H, 1), 9)) if want_all else None
        flags = C.c_uint32(0)
        rc = self.lib.sfmba_fund_ransac_host(
            self.ctx, n, H, _ptr(pairs), _ptr(samples), float(inlier_threshold), _ptr(F), _ptr(best), _ptr(mask),
            None if counts is None else _ptr(counts), None if models is None else _ptr(models), C.byref(flags))
        _capi.check(rc, self.ctx)
        out = dict(hyp=int(best[0]), count=int(best[1]), mask=mask[:n].astype(bool), F=F.reshape(3, 3),
                   flags=int(flags.value))
        if want_all:
            out.update(counts=counts[:H], models=models[:H])
        return out

    def build_visibility(self, diag_rows, num_pts: int):
        """a3/a4 on the GPU: returns int32 (obs_pt, obs_cam, obs_key, rowptr), bit-exact with
        KeyTracker.is_visible in the loop order of ba_processor.py:304-310."""
        import torch
        dev = torch.device('cuda', self.device)
        rows = [np.ascontiguousarray(r, dtype=np.int64).reshape(-1) for r in diag_rows]
        off = np.zeros(len(rows) + 1, dtype=np.int64)
        off[1:] = np.cumsum([r.shape[0] for r in rows])
        total = int(off[-1])
        diag = torch.from_numpy(np.concatenate(rows) if total else np.zeros(1, dtype=np.int64)).to(dev)
        cap = max(total, 1)
        obs_pt = torch.empty(cap, dtype=torch.int32, device=dev)
        obs_cam = torch.empty(cap, dtype=torch.int32, device=dev)
        obs_key = torch.empty(cap, dtype=torch.int32, device=dev)
        rowptr = torch.empty(num_pts + 1, dtype=torch.int32, device=dev)
        m = C.c_int32(0)
        st = torch.cuda.current_stream(self.device).cuda_stream
        rc = self.lib.sfmba_build_visibility(self.ctx, len(rows), num_pts, diag.data_ptr(), _ptr(off),
                                             obs_pt.data_ptr(), obs_cam.data_ptr(), obs_key.data_ptr(),
                                             rowptr.data_ptr(), C.byref(m), C.c_void_p(st))
        _capi.check(rc, self.ctx)
        n = m.value
        return (obs_pt[:n].cpu().numpy(), obs_cam[:n].cpu().numpy(), obs_key[:n].cpu().numpy(),
                rowptr.cpu().numpy())

    # -- profiling ---------------------------------------------------------------------------------
    def profile_enable(self, on=True, phases=None):
        """on: per-kernel spans (serialises the launches); phases: the per-phase times of SolveResult.ms (default: as on)."""
        ph = on if phases is None else phases
        self.lib.sfmba_profile_enable(self.ctx, (1 if on else 0) | (2 if ph else 0))

    def profile_reset(self):
        self.lib.sfmba_profile_reset(self.ctx)

    def profile_get(self, name: str):
        ms = C.c_double(0.0)
        n = C.c_int64(0)
        self.lib.sfmba_profile_get(self.ctx, name.encode(), C.byref(ms), C.byref(n))
        return ms.value, n.value

    def total_launches(self) -> int:
        return int(self.lib.sfmba_total_launches(self.ctx))


_default_engine = None


def get_engine() -> Engine:
    """Process-wide engine on cuda:0 (or LOCAL_RANK), created on first use."""
    global _default_engine
    if _default_engine is None:
        import os
        _default_engine = Engine(device=int(os.environ.get('LOCAL_RANK', '0')))
    return _default_engine
