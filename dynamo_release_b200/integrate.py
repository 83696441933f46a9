"""Batched integration of initial states through the RKHS vector field.

Drop-in for reference prediction/utils.py:62-259 (`integrate_vf_ivp`) and BaseVectorField.integrate
(vectorfield/scVectorField.py:721-784).  The reference runs scipy.integrate.solve_ivp once per cell, every
right-hand-side call being a NumPy `vector_field_function`; here every trajectory is one CTA of K11
(csrc/kern_ivp.cu, dvf_integrate_rk45) that runs the same adaptive Dormand-Prince 5(4) integration, terminal event
and dense output on the device.  The sampling of the finished trajectories (arc length / log space / uniform indices)
is the reference's host-side bookkeeping, restated below with the same NumPy / SciPy calls it makes.

The field must be this package's RKHS field: an object carrying `vf_dict` (SvcVectorField.func), or the plain
function dynamo itself builds around vector_field_function (`lambda x: scale * vector_field_function(x, VecFld, dims)`,
prediction/fate.py:133-135), which is recognised by tracing it (_field_of).  There is no CPU fallback for arbitrary
Python callables.
"""
from __future__ import annotations

import ctypes as C

import os

import numpy as np
import scipy.fft
import torch

from . import _lib
from .device import dev, host, lib, ptr, stream_ptr
from .vf import RkhsField, _vf_parts

RTOL, ATOL, REST_TOL = 1e-3, 1e-6, 1e-5     # solve_ivp defaults; event threshold of prediction/utils.py:107
_BYTES_PER_LAUNCH = 1 << 30                 # trajectories are integrated in groups whose outputs fit this
PROFILE = False                             # developer aid (tools/ivp_tune.py): CUDA-event time of the kernel launches
KERNEL_MS = 0.0
_FFT_WORKERS = max(1, min(32, os.cpu_count() or 1))


class ScaledRkhsField(RkhsField):
    """x -> scale * f(x): what BaseVectorField.integrate binds when `func` is unset (scVectorField.py:764-766)."""

    def __init__(self, vf_dict, scale=1.0):
        super().__init__(vf_dict)
        self.scale = scale

    def __call__(self, x, **kw):
        return self.scale * super().__call__(x, **kw)


class _DeviceBatch:
    """What one dvf_integrate_rk45 launch left in HBM (times, states, interpolant coefficients of a group of
    trajectories); kept alive by the solutions that refer to it so that they can be evaluated on the device."""

    def __init__(self, ts, ys, Q, ns, st, D, cap):
        self.ts, self.ys, self.Q, self.ns, self.st, self.D, self.cap = ts, ys, Q, ns, st, D, cap


class DenseSolution:
    """The continuous solution of one trajectory.  Called with times it evaluates like scipy's OdeSolution over
    RkDenseOutput pieces (scipy/integrate/_ivp/common.py, rk.py), for all requested times at once, in NumPy; solutions
    produced by solve_ivp_batch keep their coefficients on the device (fetched only if this NumPy path is used) and
    are evaluated there, many trajectories per launch, by dense_eval_many()."""

    def __init__(self, ts, ys, Q=None, h=None, batch=None, slot=None):
        ts = np.asarray(ts)
        d = np.diff(ts)
        if not ((ts.size == 2 and ts[0] == ts[-1]) or np.all(d > 0) or np.all(d < 0)):
            raise ValueError("`ts` must be strictly increasing or decreasing.")
        self.ts, self.ys, self._Q, self.h = ts, ys, Q, (d if h is None else np.asarray(h))
        self.batch, self.slot = batch, slot
        self.n_segments = len(d)
        self.ascending = bool(ts[-1] >= ts[0])
        self.t_min, self.t_max = (ts[0], ts[-1]) if self.ascending else (ts[-1], ts[0])

    @property
    def Q(self):
        if self._Q is None:
            self._Q = self.batch.Q[self.slot, :self.n_segments].cpu().numpy()
        return self._Q

    def __call__(self, t):
        t = np.asarray(t, dtype=np.float64)
        tt = np.atleast_1d(t)
        if self.ascending:
            seg = np.searchsorted(self.ts, tt, side="left") - 1
            seg = np.clip(seg, 0, self.n_segments - 1)
        else:
            seg = np.searchsorted(self.ts[::-1], tt, side="right") - 1
            seg = self.n_segments - 1 - np.clip(seg, 0, self.n_segments - 1)
        h = self.h[seg]
        x = (tt - self.ts[seg]) / h
        p = np.cumprod(np.tile(x, (4, 1)), axis=0)                       # x, x^2, x^3, x^4
        Qs = self.Q[seg]                                                  # (nt, D, 4)
        y = ((Qs[:, :, 0] * p[0][:, None] + Qs[:, :, 1] * p[1][:, None]) + Qs[:, :, 2] * p[2][:, None]) \
            + Qs[:, :, 3] * p[3][:, None]
        y = (h[:, None] * y).T + self.ys[:, seg]
        return y[:, 0] if t.ndim == 0 else y


def dense_eval_many(sols, times):
    """[sol(t) for sol, t in zip(sols, times)] -> list of (D, len(t)) arrays, with one dvf_ivp_dense_eval launch per
    group of device-resident solutions (the interpolant coefficients never leave HBM)."""
    out = [None] * len(sols)
    groups = {}
    for r, sol in enumerate(sols):
        if sol.batch is None:
            out[r] = sol(np.asarray(times[r], dtype=np.float64))        # built from host arrays by the caller
        else:
            groups.setdefault(id(sol.batch), (sol.batch, []))[1].append(r)
    for batch, rows in groups.values():
        m, D = len(rows), batch.D
        ne = np.array([len(times[r]) for r in rows], dtype=np.int32)
        S = int(ne.max())
        if S == 0:
            for r in rows:
                out[r] = np.empty((D, 0))
            continue
        te = np.zeros((m, S))
        for k, r in enumerate(rows):
            te[k, :ne[k]] = times[r]
        traj = np.array([sols[r].slot for r in rows], dtype=np.int32)
        te_d, ne_d, traj_d = dev(te), dev(ne, torch.int32), dev(traj, torch.int32)
        res = torch.empty((m, D, S), dtype=torch.float64, device="cuda")
        _lib.check(lib().dvf_ivp_dense_eval(ptr(batch.ts), ptr(batch.ys), ptr(batch.Q), ptr(batch.ns), ptr(batch.st),
                                            D, batch.cap, ptr(traj_d), ptr(te_d), ptr(ne_d), m, S, ptr(res),
                                            stream_ptr()))
        res_h = host(res)
        for k, r in enumerate(rows):
            out[r] = np.ascontiguousarray(res_h[k, :, :ne[k]])
    return out


class IvpResult:
    """t (n+1,), y (D, n+1), sol, status: the fields of scipy's solve_ivp result the reference reads (:124,135)."""

    def __init__(self, t, y, sol, status):
        self.t, self.y, self.sol, self.status = t, y, sol, status


def _exact_scale(s, pairs):
    """The factor the callable multiplied the field by: the float next to the ratio s for which factor * field
    reproduces the callable's outputs bit for bit (the ratio of two rounded numbers can be one ulp off)."""
    cands = [s, np.nextafter(s, np.inf), np.nextafter(s, -np.inf)] + [float(np.round(s, k)) for k in range(16)]
    for c in cands:
        if all(np.array_equal(c * r, o) for r, o in pairs):
            return float(c)
    return float(s)


def _field_of(f, probes=None):
    """(vf_dict, scale) of the field behind the callable `f`.  Objects carrying `vf_dict` (RkhsField, what
    SvcVectorField.func / vecfld_from_adata return) are taken as they are.  A plain function -- dynamo builds
    `lambda x: scale * vector_field_function(x=x, vf_dict=VecFld, dim=dims)` in prediction/fate.py:133-135 and
    scVectorField.py:764-766 -- is called on a few of the initial states with a trace open on vector_field_function
    (vf._TRACE): if every call evaluated one and the same RKHS field over all its dimensions AT THE POINT IT WAS GIVEN
    (a wrapper that rescales or shifts its argument first, `lambda x: vff((x - xm) / xscale, vf)`, is not the field)
    and the output is one constant times that value, the callable IS scale x field and runs on the device.  At least
    three distinct points are probed (perturbed copies of the states are added when fewer are given), so a factor that
    depends on the state (`lambda x: vff(x, vf) * x[0]`) is seen to vary.  Anything else is refused: there is no CPU
    path for arbitrary Python callables."""
    from . import vf as _vf

    vf_dict = getattr(f, "vf_dict", None)
    if vf_dict is not None:
        return vf_dict, float(getattr(f, "scale", 1.0))
    refuse = NotImplementedError(
        "integrate_vf_ivp on the B200 backend integrates this package's RKHS field: pass an object carrying vf_dict "
        "(SvcVectorField.func / RkhsField) or a function of the form `lambda x: scale * vector_field_function(x, vf_dict)`; "
        "other Python callables have no device path and there is no CPU fallback")
    if probes is None or len(probes) == 0 or not callable(f):
        raise refuse
    pairs, fields = [], []
    pts = [np.array(x, dtype=np.float64) for x in np.atleast_2d(probes)[:3]]
    k = 0
    while len(pts) < 3:                     # perturbed copies: distinct points even for a single initial state
        base = pts[k % len(pts)]
        k += 1
        pts.append(base * (1.0 + 0.03125 * k) + 0.0078125 * k * (1.0 + np.abs(base).max()))
    for x in pts:
        _vf._TRACE.rec = rec = []
        try:
            out = np.asarray(f(np.array(x, dtype=np.float64)))
        finally:
            _vf._TRACE.rec = None
        if len(rec) != 1:
            raise refuse
        x_seen, d, dim, ctrl_ind, res = rec[0]
        if x_seen.shape != x.shape or not np.array_equal(x_seen, x):
            raise refuse                        # the callable evaluates the field somewhere else than at its argument
        res = np.asarray(res)
        n_dim = np.shape(d["X_ctrl"])[1]
        full = dim is None or (np.isscalar(dim) and dim == n_dim) or \
            (not np.isscalar(dim) and list(np.atleast_1d(dim)) == list(range(n_dim)))
        if ctrl_ind is not None or not full or out.shape != res.shape:
            raise refuse
        fields.append(d)
        pairs.append((res, out))
    if any(d is not fields[0] for d in fields):
        raise refuse
    ratios = [o.flat[np.argmax(np.abs(r))] / r.flat[np.argmax(np.abs(r))] for r, o in pairs if np.any(r != 0)]
    if not ratios:
        if all(not np.any(o) for _, o in pairs):
            return fields[0], 1.0              # the field vanishes at the probes and so does the callable
        raise refuse
    s = ratios[0]
    for r, o in pairs:
        if not np.allclose(o, s * r, rtol=1e-12, atol=0.0):
            raise refuse                        # not a constant multiple of the field
    return fields[0], _exact_scale(s, pairs)


def solve_ivp_batch(init_states, vf_dict, t_span, max_step, scale=1.0, rtol=RTOL, atol=ATOL, cap=None):
    """All initial states (n, D) from t_span[0] to t_span[1] on the device -> list of IvpResult, one per state,
    equal to `solve_ivp(lambda t, x: scale * f(x), t_span, y0, events=<all |f| < 1e-5, terminal>, max_step=max_step,
    dense_output=True)` of reference prediction/utils.py:115-135."""
    Xc, beta, Cm = _vf_parts(vf_dict)
    y0 = dev(np.ascontiguousarray(np.atleast_2d(np.asarray(init_states, dtype=np.float64))))
    Xc, Cm = dev(Xc), dev(Cm)
    n, D = y0.shape
    M = Xc.shape[0]
    if Xc.shape[1] != D:
        raise ValueError("init_states must have the %d columns of X_ctrl, got %d" % (Xc.shape[1], D))
    t0, tb = float(t_span[0]), float(t_span[1])
    max_step = float(max_step)
    if not max_step > 0:
        raise ValueError("`max_step` must be positive.")      # scipy validate_max_step
    span = abs(tb - t0)
    if cap is None:
        cap = int(min(max(span / max_step, 1.0), 1e6) * 1.25) + 64 if np.isfinite(max_step) else 1024
    L = lib()
    nbytes = C.c_size_t(0)
    _lib.check(L.dvf_ivp_workspace_bytes(D, M, C.byref(nbytes)))
    ws = torch.empty((nbytes.value + 7) // 8, dtype=torch.float64, device="cuda")
    wave = 16 * torch.cuda.get_device_properties(y0.device).multi_processor_count
    out = [None] * n
    todo = np.arange(n)
    while todo.size:
        per = max(1, int(_BYTES_PER_LAUNCH // (8 * (cap + 1) * (5 * D + 1))))
        if per >= wave:
            per -= per % wave                 # whole waves of the lock-step kernel (4 CTAs per SM x 4 trajectories)
        again = []
        for c0 in range(0, todo.size, per):
            idx = todo[c0:c0 + per]
            k = idx.size
            ysel = y0[torch.as_tensor(idx, device=y0.device)].contiguous() if k != n else y0
            ts = torch.empty((k, cap + 2), dtype=torch.float64, device="cuda")
            ys = torch.empty((k, cap + 1, D), dtype=torch.float64, device="cuda")
            Q = torch.empty((k, cap, D, 4), dtype=torch.float64, device="cuda")
            ns = torch.empty((k,), dtype=torch.int32, device="cuda")
            st = torch.empty((k,), dtype=torch.int32, device="cuda")
            if PROFILE:
                ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                ev0.record()
            _lib.check(L.dvf_integrate_rk45(ptr(ysel), k, D, ptr(Xc), M, float(beta), ptr(Cm), float(scale), t0, tb,
                                            max_step, float(rtol), float(atol), REST_TOL, cap, ptr(ts), ptr(ys),
                                            ptr(Q), ptr(ns), ptr(st), ptr(ws), nbytes.value, stream_ptr()))
            if PROFILE:
                global KERNEL_MS
                ev1.record()
                ev1.synchronize()
                KERNEL_MS += ev0.elapsed_time(ev1)
            ns_h, st_h = ns.cpu().numpy(), st.cpu().numpy()
            kmax = int(ns_h.max()) if k else 0
            ts_h = host(ts[:, :kmax + 2].contiguous())            # times and states go to the host (the reference's
            ys_h = host(ys[:, :kmax + 1].contiguous())            # result fields); Q stays in HBM
            batch = _DeviceBatch(ts, ys, Q, ns, st, D, cap)
            for j, i in enumerate(idx):
                if st_h[j] == -2:
                    again.append(i)
                    continue
                m = int(ns_h[j])
                t_j = ts_h[j, :m + 1]            # views of the group's pinned host buffers
                y_j = ys_h[j, :m + 1].T
                h_j = None
                status = int(st_h[j])
                if status == -3:
                    # the event's root is the START of the first step (initial state and step end both at rest):
                    # scipy returns t = [t0, t0], status 1, and an interpolant that keeps the step as taken
                    h_j, status = np.array([ts_h[j, 2] - ts_h[j, 0]]), 1
                out[i] = IvpResult(t_j, y_j, DenseSolution(t_j, y_j, None, h_j, batch=batch, slot=j), status)
        todo = np.asarray(again, dtype=np.int64)
        cap *= 2
    return out


# ---- the reference's trajectory bookkeeping (host, O(steps x D) per trajectory) --------------------------------
def _half_spectra(a):
    """vectorfield/topography.py:324-329,349-354 (calc_fft per column: amplitudes without the mean term, columns
    concatenated) for a stack of trajectories a (k, n, D) -> (k, D * (n // 2 - 1)): one transform along the time axis
    for all trajectories and columns -- the same numbers as the reference's per-column calls."""
    n = a.shape[1]
    # scipy.fft runs the same pocketfft transforms as numpy.fft (bit-identical, tests/test_ivp_cpu.py) on all host cores
    amp = (np.abs(scipy.fft.rfft(a, axis=1, workers=_FFT_WORKERS)) / n * 2)[:, : int(n / 2)][:, 1:]
    return np.ascontiguousarray(amp.transpose(0, 2, 1)).reshape(len(a), -1)


def _repeat_cuts(xs, max_iter=5, n_dom=3):
    """vectorfield/topography.py:332-397 (dup_osc_idx_iter over dup_osc_idx) for many trajectories: as long as the
    spectra of the last two thirds of x[:idx] agree within tol, idx shrinks to the end of the second third.  The
    sequence of candidate cuts depends on the length only, so trajectories of equal length go through the rounds
    together (one batched FFT per round); tol as in prediction/trajectory.py:155."""
    out = np.array([len(x) for x in xs], dtype=np.int64)
    by_len = {}
    for i, x in enumerate(xs):
        by_len.setdefault(len(x), []).append(i)
    for length, members in by_len.items():
        act = np.array(members)
        stack = np.stack([xs[i] for i in members])                          # (k, length, D)
        tol = (stack.max(1) - stack.min(1)).mean(1) / 1000
        idx, j = length, 0
        while act.size:
            L = int(np.floor(idx / n_dom))
            try:
                sa = _half_spectra(stack[:, (n_dom - 2) * L:(n_dom - 1) * L])
                sb = _half_spectra(stack[:, (n_dom - 1) * L:n_dom * L])
            except ValueError:          # empty thirds: the reference reports "calc_fft_k run failed" and stops
                break
            with np.errstate(invalid="ignore", divide="ignore"):
                d = sa - sb
                diff = np.sqrt(np.einsum("ij,ij->i", d, d)) / sa.shape[1]
                # the reference's np.linalg.norm sums in BLAS order: decide with it wherever the two could differ
                for r in np.flatnonzero(np.abs(diff - tol) <= 1e-9 * tol):
                    diff[r] = np.linalg.norm(d[r]) / sa.shape[1]
            keep = diff <= tol
            idx = (n_dom - 1) * L
            if not keep.all():
                act, stack, tol = act[keep], stack[keep], tol[keep]
            out[act] = idx
            j += 1
            if j >= max_iter or idx == 0:
                break
    return out


def _arclength_times(x, num, t):
    """prediction/trajectory.py:568-592 (arclength_sampling_n), the resampled times only.  Its
    interp1d(arclen, t)(z) is, for these 1-d float64 arrays, scipy's delegation to numpy.interp."""
    arclen = np.hstack((0, np.cumsum(np.linalg.norm(np.diff(x, axis=0), axis=1))))
    if len(arclen) < 2:
        raise ValueError("x and y arrays must have at least 2 entries")     # interp1d's own check
    z = np.linspace(arclen[0], arclen[-1], num)
    return np.interp(z, arclen, t)


def _eval_all(SOL, times, both):
    """[SOL[i](times[i])] for every trajectory (prediction/trajectory.py:167-176, prediction/utils.py:231-242): for
    "both" the negative times go to the backward solution and the rest to the forward one."""
    if not both:
        return dense_eval_many(SOL, times)
    neg = [int(np.sum(np.asarray(t) < 0)) for t in times]
    back = dense_eval_many([s[0] for s in SOL], [t[:k] for t, k in zip(times, neg)])
    fwd = dense_eval_many([s[1] for s in SOL], [t[k:] for t, k in zip(times, neg)])
    return [np.hstack((b, f)) for b, f in zip(back, fwd)]


RESIDENT_BYTES = None      # HBM budget for the dense solutions of one slice of cells; None = half of the free memory


def _resident_budget():
    if RESIDENT_BYTES is not None:
        return RESIDENT_BYTES
    if torch.cuda.is_available():
        return torch.cuda.mem_get_info()[0] // 2
    return 1 << 34


def _resident_bytes_per_cell(t, interpolation_num, D):
    """What solve_ivp_batch keeps on the device per trajectory (its step capacity x times, states, coefficients)."""
    cap = int(min(max(float(interpolation_num), 1.0), 1e6) * 1.25) + 64
    return 8 * (cap + 2) * (5 * D + 1)


def integrate_vf_ivp(init_states, t, integration_direction, f, args=None, interpolation_num=250, average=True,
                     sampling="arc_length", verbose=False, disable=False):
    """reference prediction/utils.py:62-259.  Returns (t, Y): lists with one time vector and one (D, n_points) array
    per initial state."""
    init_states = np.asarray(init_states)
    if init_states.ndim == 1:
        init_states = init_states[None, :]
    n_cell, n_feature = init_states.shape
    if n_cell == 0:
        return [], []                       # the reference's per-cell loops do not run
    vf_dict, scale = _field_of(f, init_states)
    f = ScaledRkhsField(vf_dict, scale)     # what the callable was recognised as (slices below do not trace again)
    t = np.asarray(t)
    if integration_direction not in ("forward", "backward", "both"):
        raise Exception("both, forward, backward are the only valid direction argument strings")
    # The dense solutions stay in HBM until they have been resampled.  arc_length / logspace treat every trajectory on
    # its own, so a large set of cells goes through solve -> resample -> evaluate in slices that fit the memory budget
    # (uniform_indices shares one time grid between all trajectories and needs them all at once).
    per_cell = _resident_bytes_per_cell(t, interpolation_num, n_feature) * (2 if integration_direction == "both" else 1)
    slice_cells = max(1, int(_resident_budget() // per_cell))
    if sampling in ("arc_length", "logspace") and n_cell > slice_cells:
        tt, YY = [], []
        for c0 in range(0, n_cell, slice_cells):
            a, b = integrate_vf_ivp(init_states[c0:c0 + slice_cells], t, integration_direction, f, args,
                                    interpolation_num, average, sampling, verbose, disable)
            tt += a
            YY += b
        return tt, YY
    max_step = np.abs(t[-1] - t[0]) / interpolation_num                      # :96
    if interpolation_num is not None and integration_direction == "both":
        interpolation_num = interpolation_num * 2                             # :100-101
    both = integration_direction == "both"
    if integration_direction == "forward":
        res = solve_ivp_batch(init_states, vf_dict, (t[0], t[-1]), max_step, scale)
        T, Y, SOL = [r.t for r in res], [r.y for r in res], [r.sol for r in res]
    elif integration_direction == "backward":
        res = solve_ivp_batch(init_states, vf_dict, (-t[0], -t[-1]), max_step, scale)
        T, Y, SOL = [r.t for r in res], [r.y for r in res], [r.sol for r in res]
    elif both:
        fw = solve_ivp_batch(init_states, vf_dict, (t[0], t[-1]), max_step, scale)
        bw = solve_ivp_batch(init_states, vf_dict, (-t[0], -t[-1]), max_step, scale)
        # :155-158 as written there: the backward STATES are flipped along the feature axis, the times along time
        Y = [np.hstack((b.y[::-1, :], a.y)) for a, b in zip(fw, bw)]
        T = [np.hstack((b.t[::-1], a.t)) for a, b in zip(fw, bw)]
        SOL = [[b.sol, a.sol] for a, b in zip(fw, bw)]
    else:
        raise Exception("both, forward, backward are the only valid direction argument strings")

    if sampling == "arc_length":                                             # :172-184, trajectory.py:141-176
        cuts = _repeat_cuts([y.T for y in Y], max_iter=100)
        t_out = [_arclength_times(Y[i].T[:cuts[i]], interpolation_num + 1, T[i][:cuts[i]])[1:] for i in range(n_cell)]
        return t_out, _eval_all(SOL, t_out, both)
    if sampling == "logspace":                                               # :185-197, trajectory.py:178-224
        t_out = []
        for i in range(n_cell):
            tau = T[i]
            neg_tau, pos_tau = tau[tau < 0], tau[tau >= 0]
            if len(neg_tau) > 0:
                t_0 = -(np.logspace(0, np.log10(abs(min(neg_tau)) + 1), interpolation_num)) - 1
                t_1 = np.logspace(0, np.log10(max(pos_tau) + 1), interpolation_num) - 1
                ti = np.hstack((t_0[::-1], t_1))
            else:
                ti = np.logspace(0, np.log10(max(tau) + 1), interpolation_num) - 1
            t_out.append(ti)
        return t_out, _eval_all(SOL, t_out, both)
    if sampling == "uniform_indices":                                        # :198-257
        t_uniq = np.unique(np.hstack(T))
        if len(t_uniq) > interpolation_num:
            valid_t = np.hstack([0, np.sort(t_uniq)])[(np.linspace(0, len(t_uniq), interpolation_num)).astype(int)]
            if len(valid_t) != interpolation_num:
                n_missed = interpolation_num - len(valid_t)
                tmp = np.zeros(n_missed)
                for i in range(n_missed):
                    tmp[i] = (valid_t[i] + valid_t[i + 1]) / 2
                valid_t = np.sort(np.hstack([tmp, valid_t]))
        else:
            valid_t = np.hstack((-np.linspace(min(t_uniq), 0, interpolation_num),
                                 np.linspace(0, max(t_uniq), interpolation_num)))
        Yall = np.hstack(_eval_all(SOL, [valid_t] * n_cell, both))
        tt = valid_t
        if n_cell > 1 and average:
            t_len = int(len(tt) / n_cell)
            avg = np.zeros((n_feature, t_len))
            for i in range(t_len):
                avg[:, i] = np.mean(Yall[:, np.arange(n_cell) * t_len + i], 1)
            Yall = avg
        width = Yall.shape[1] // n_cell
        return [tt] * n_cell, [Yall[:, i * width:(i + 1) * width] for i in range(n_cell)]
    # the reference falls through with the raw solver output when `sampling` is none of the three
    return t, Y


def _fate(VecFld, init_states, t_end=None, step_size=None, direction="both", interpolation_num=250, average=True,
          sampling="arc_length", cores=1):
    """reference prediction/fate.py:166-265: the time sequence of getTseq, integrate_vf_ivp, and -- for several cells
    with `average` -- the mean state per sampled time point.  `cores` is accepted and ignored: the reference's thread
    pool calls integrate_vf_ivp once per cell, which for these per-trajectory samplings gives what one batched launch
    gives (it forces cores = 1 itself for "uniform_indices")."""
    from .svc import _t_seq

    init_states = np.asarray(init_states)
    t, prediction = integrate_vf_ivp(init_states, _t_seq(init_states, t_end, step_size), direction, VecFld,
                                     interpolation_num=interpolation_num, average=average, sampling=sampling)
    if init_states.shape[0] > 1 and average:                                  # :246-262
        t_stack, prediction_stack = np.hstack(t), np.hstack(prediction)
        n_cell, n_feature = init_states.shape
        if len(t_stack) > 0 and len(prediction_stack) > 0:
            t_len = int(len(t_stack) / n_cell)
            avg = np.zeros((n_feature, t_len))
            for i in range(t_len):
                avg[:, i] = np.mean(prediction_stack[:, np.arange(n_cell) * t_len + i], 1)
            prediction = [avg]
            t = [np.sort(np.unique(t))]
    return t, prediction
