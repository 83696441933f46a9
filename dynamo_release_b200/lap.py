"""Least-action-path functional on the B200 field: drop-ins for dynamo.vectorfield.least_action_path (reference
vectorfield/least_action_path.py:7-120) -- `action`, `action_aux`, `action_grad`, `action_grad_aux`, `reshape_path`,
`least_action_path` with the reference's signatures -- plus batched entries for many paths at once.

The reference evaluates f at the segment centres, builds every Jacobian (D, D, P-1) and multiplies row by row on the
host (:46-52).  Here one launch (dvf_lap_terms, kern_eval.cu) returns dv = v - f(x) and z = dv J(x) for all segments of
all paths without materialising J.  The device path needs to know the field: `vf_func` / `jac_func` must be this
package's RkhsField / RkhsJacobian (what SvcVectorField.func / get_Jacobian() return) over the same VecFldDict; other
callables raise NotImplementedError (there is no CPU fallback).  `least_action_path` keeps scipy's `minimize` driving
the iterations on the host, as the reference does (:115).
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib
from .device import dev, empty, host, lib, ptr, stream_ptr
from .vf import RkhsField, RkhsJacobian, _vf_parts


def _field_dict(vf_func, jac_func=None):
    d = getattr(vf_func, "vf_dict", None)
    if d is None or (jac_func is not None and getattr(jac_func, "vf_dict", None) is not d):
        raise NotImplementedError(
            "the least-action functional on the B200 backend needs this package's field objects over ONE VecFldDict "
            "(vf_func = SvcVectorField.func / RkhsField, jac_func = SvcVectorField.get_Jacobian() / RkhsJacobian); "
            "other Python callables have no device path and there is no CPU fallback")
    return d


def lap_terms(x, vp, vf_dict, want_z=True, device_out=False):
    """dv = vp - f(x) and z = dv J(x) for the rows of x (n, D): one launch."""
    Xc, beta, Cm = _vf_parts(vf_dict)
    x, vp, Xc, Cm = dev(x), dev(vp), dev(Xc), dev(Cm)
    n, D = x.shape
    dv = empty((n, D))
    z = empty((n, D)) if want_z else None
    _lib.check(lib().dvf_lap_terms(ptr(x), ptr(vp), n, D, ptr(Xc), Xc.shape[0], float(beta), ptr(Cm), ptr(dv), ptr(z),
                                   stream_ptr()))
    if device_out:
        return dv, z
    return host(dv), (host(z) if want_z else None)


def action(path, vf_func, D=1, dt=1):
    """reference least_action_path.py:7-27."""
    vf_dict = _field_dict(vf_func)
    path = np.asarray(path, dtype=np.float64)
    x = (path[:-1] + path[1:]) * 0.5
    v = np.diff(path, axis=0) / dt
    dv, _ = lap_terms(x, v, vf_dict, want_z=False)
    s = dv.flatten()
    return 0.5 * s.dot(s) * dt / D


def action_aux(path_flatten, vf_func, dim, start=None, end=None, **kwargs):
    """reference :30-32."""
    return action(reshape_path(path_flatten, dim, start=start, end=end), vf_func, **kwargs)


def action_grad(path, vf_func, jac_func, D=1, dt=1):
    """reference :35-57: gradient of the action with respect to the interior points of the path."""
    vf_dict = _field_dict(vf_func, jac_func)
    path = np.asarray(path, dtype=np.float64)
    x = (path[:-1] + path[1:]) * 0.5
    v = np.diff(path, axis=0) / dt
    dv, z = lap_terms(x, v, vf_dict)
    return (dv[:-1] - dv[1:]) / D - dt / (2 * D) * (z[:-1] + z[1:])


def action_grad_aux(path_flatten, vf_func, jac_func, dim, start=None, end=None, **kwargs):
    """reference :60-62."""
    return action_grad(reshape_path(path_flatten, dim, start=start, end=end), vf_func, jac_func, **kwargs).flatten()


def reshape_path(path_flatten, dim, start=None, end=None):
    """reference :65-71."""
    path = path_flatten.reshape(int(len(path_flatten) / dim), dim)
    if start is not None:
        path = np.vstack((start, path))
    if end is not None:
        path = np.vstack((path, end))
    return path


def least_action_path(start, end, vf_func, jac_func, n_points=20, init_path=None, D=1):
    """reference :74-120: scipy's minimize over the interior points, action and gradient evaluated on the device."""
    from scipy.optimize import minimize

    dim = len(start)
    if init_path is None:
        path_0 = (
            np.tile(start, (n_points + 1, 1))
            + (np.linspace(0, 1, n_points + 1, endpoint=True) * np.tile(end - start, (n_points + 1, 1)).T).T
        )
    else:
        path_0 = init_path
    fun = lambda x: action_aux(x, vf_func, dim, start=path_0[0], end=path_0[-1], D=D)  # noqa: E731
    jac = lambda x: action_grad_aux(x, vf_func, jac_func, dim, start=path_0[0], end=path_0[-1], D=D)  # noqa: E731
    # the reference passes the 2-D block path_0[1:-1] (:115), which scipy >= 1.11 refuses ("'x0' must only have one
    # dimension"); flattened as prediction/least_action_path.py:388 does
    sol_dict = minimize(fun, np.asarray(path_0[1:-1]).flatten(), jac=jac)
    path_sol = reshape_path(sol_dict["x"], dim, start=path_0[0], end=path_0[-1])
    return path_sol, sol_dict


def action_and_grad_batch(paths, vf_dict, D=1, dt=1):
    """Many paths at once (B, P, D_state): actions (B,) and gradients (B, P-2, D_state) from ONE launch over all
    B (P-1) segments -- what a search over many cell pairs (prediction/least_action_path.py:300-360 loops over them)
    evaluates per iteration."""
    paths = np.asarray(paths, dtype=np.float64)
    B, P, Ds = paths.shape
    x = ((paths[:, :-1] + paths[:, 1:]) * 0.5).reshape(B * (P - 1), Ds)
    v = (np.diff(paths, axis=1) / dt).reshape(B * (P - 1), Ds)
    dv, z = lap_terms(x, v, vf_dict)
    dv, z = dv.reshape(B, P - 1, Ds), z.reshape(B, P - 1, Ds)
    act = 0.5 * np.einsum("bpd,bpd->b", dv, dv) * dt / D
    grad = (dv[:, :-1] - dv[:, 1:]) / D - dt / (2 * D) * (z[:, :-1] + z[:, 1:])
    return act, grad
