"""SparseVFC on the GPU: drop-in for dynamo.vectorfield.scVectorField.SparseVFC (reference :313-542).

Same positional signature, defaults, exceptions and returned VecFldDict layout (vectorfield/utils.py:34-51).
The prelude (prelude.py) fixes the control points; the EM loop runs inside one C-ABI call (dvf_fit_em).
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib, kernels, prelude
from .device import dev, empty, host, lib, ptr, stream_ptr


_POOL = None


def _copy_pool():
    global _POOL
    if _POOL is None:
        from concurrent.futures import ThreadPoolExecutor
        _POOL = ThreadPoolExecutor(max_workers=2, thread_name_prefix="dvf-host-copy")
    return _POOL


class FitState:
    """Device-resident result of a fit (benchmark mode: nothing is copied back unless asked)."""

    def __init__(self):
        self.C = self.P = self.V = self.E_traj = self.tecr_traj = None
        self.n_iter = 0
        self.sigma2 = self.gamma = self.n_inlier = float("nan")
        self.workspace = None


def solver_code(lstsq_method):
    """`lstsq_method` of SparseVFC (scVectorField.py:185-194) -> DVF_SOLVER_*.

    "drouin" (the reference's default; also what the reference falls back to for an unknown name, :191-193): the
    reference's own form, normal equations solve(lhs^T lhs, lhs^T rhs) by LU with partial pivoting, on the device.
    "cholesky": an addition of this backend -- factor the symmetric positive definite lhs itself (error cond eps
    instead of cond^2 eps; what round 1 always did).  "scipy" (scipy.linalg.lstsq, an SVD) is not built."""
    if lstsq_method == "cholesky":
        return _lib.SOLVER_CHOLESKY
    if lstsq_method == "scipy":
        raise NotImplementedError("lstsq_method='scipy' (scipy.linalg.lstsq / SVD) is not served by the B200 backend; "
                                  "use 'drouin' (the reference's default) or 'cholesky'")
    return _lib.SOLVER_NORMAL_LU


def fit_em(X, Y, Xc, beta, a=5, gamma=0.9, lambda_=3, minP=1e-5, ecr=1e-5, theta=0.75, MaxIter=500,
           comm=None, N_total=None, check_every=0, workspace=None, solver=_lib.SOLVER_NORMAL_LU) -> FitState:
    """EM loop of scVectorField.py:417-497 on device tensors X, Y (N x D, valid rows only), Xc (M x D)."""
    X, Y, Xc = dev(X), dev(Y), dev(Xc)
    N, D = X.shape
    M = Xc.shape[0]
    if D > kernels.MAX_D:
        raise NotImplementedError("D=%d > %d is not supported by the sm_100a kernels yet" % (D, kernels.MAX_D))
    L = lib()
    nbytes = C.c_size_t(0)
    _lib.check(L.dvf_fit_workspace_bytes(N, D, M, C.byref(nbytes)))
    if workspace is None or workspace.numel() * 8 < nbytes.value:
        workspace = torch.empty((nbytes.value + 7) // 8, dtype=torch.float64, device="cuda")
    st = FitState()
    st.workspace = workspace
    st.C = empty((M, D))
    st.P = empty((N + 2,))
    st.V = empty((N, D))
    n_traj = max(int(MaxIter), 1)
    st.E_traj = torch.full((n_traj,), float("nan"), dtype=torch.float64, device="cuda")
    st.tecr_traj = torch.full((n_traj,), float("nan"), dtype=torch.float64, device="cuda")
    prm = _lib.EmParams(float(a), float(gamma), float(lambda_), float(minP), float(ecr), float(theta),
                        int(MaxIter), int(check_every), int(solver), 0)
    res = _lib.EmResult()
    _lib.check(L.dvf_fit_em(ptr(X), ptr(Y), N, int(N if N_total is None else N_total), D, ptr(Xc), M, float(beta),
                            C.byref(prm), ptr(st.C), ptr(st.P), ptr(st.V), ptr(st.E_traj), ptr(st.tecr_traj),
                            C.byref(res), ptr(workspace), workspace.numel() * 8, comm, stream_ptr()))
    st.P = st.P[:N]
    st.n_iter, st.sigma2, st.gamma, st.n_inlier = res.n_iter, res.sigma2, res.gamma, res.n_inlier
    return st


def SparseVFC(X, Y, Grid, M=100, a=5, beta=None, ecr=1e-5, gamma=0.9, lambda_=3, minP=1e-5, MaxIter=500,
              theta=0.75, div_cur_free_kernels=False, velocity_based_sampling=True, sigma=0.8, eta=0.5, seed=0,
              lstsq_method="drouin", verbose=1, **_ignored):
    """Drop-in for dynamo's SparseVFC (scVectorField.py:313-333).  Returns the VecFldDict with numpy values.

    Differences, all documented in DESIGN.md: the div/curl-free kernel branch raises NotImplementedError;
    `lstsq_method`: "drouin" (default) is the reference's normal-equations form on the device, "cholesky" is this
    backend's stable alternative, "scipy" raises NotImplementedError (solver_code).  `sharded=True` (or the
    environment variable DVF_SHARDED=1) inside a torchrun job: X, Y are this rank's row block and the fit is ONE
    joint fit over all ranks (dist.SparseVFC_sharded)."""
    if div_cur_free_kernels:
        raise NotImplementedError("div_cur_free_kernels=True is outside the B200 hot path (Gaussian kernel only)")
    solver = solver_code(lstsq_method)
    sharded = _ignored.pop("sharded", None)
    if sharded is None:
        import os
        sharded = os.environ.get("DVF_SHARDED", "0") not in ("", "0")
    if sharded:
        from . import dist as _dist
        if _dist.active():
            return _dist.SparseVFC_sharded(X, Y, Grid, M=M, a=a, beta=beta, ecr=ecr, gamma=gamma, lambda_=lambda_,
                                           minP=minP, MaxIter=MaxIter, theta=theta,
                                           velocity_based_sampling=velocity_based_sampling, seed=seed,
                                           lstsq_method=lstsq_method, gather=_ignored.pop("gather", True))
    # the reference copies its inputs (:396); the two host copies run on helper threads (NumPy releases the GIL)
    # while this thread uploads the data and drives the GPU
    copies = _copy_pool().map(lambda a: np.array(a, copy=True), (X, Y))
    Xd, Yd = dev(X), dev(Y)
    valid, Xv, Yv, ctrl, idx, beta = prelude.run(Xd, Yd, M, seed=seed, beta=beta,
                                                 velocity_based_sampling=velocity_based_sampling)
    N, D = Yv.shape
    if N == 0:
        raise Exception("no cell with a finite velocity vector")
    st = fit_em(Xv, Yv, ctrl, beta, a=a, gamma=gamma, lambda_=lambda_, minP=minP, ecr=ecr, theta=theta,
                MaxIter=MaxIter, solver=solver)
    if st.n_iter == 0:
        # reference :498-503 -- the loop never ran
        raise Exception(
            "please check your input parameters, "
            f"tecr: 1, ecr {ecr} and sigma2 {st.sigma2},"
            f"tecr must larger than ecr and sigma2 must larger than 1e-8"
        )
    i = st.n_iter
    X_ori, Y_ori = list(copies)
    Cm = host(st.C)
    P = host(st.P).reshape(-1, 1)
    grid_V = None
    if Grid is not None:
        grid_V = host(kernels.eval_field(Grid, ctrl, beta, st.C))  # np.dot(con_K(Grid, ctrl), C)  (:505-507)
    return {
        "X": X_ori,
        "valid_ind": host(valid),
        "X_ctrl": ctrl,
        "ctrl_idx": idx,
        "Y": Y_ori,
        "beta": beta,
        "V": host(st.V),
        "C": Cm,
        "P": P,
        "VFCIndex": np.where(P > theta)[0],
        "sigma2": st.sigma2,
        "grid": Grid,
        "grid_V": grid_V,
        "iteration": i - 1,
        "tecr_traj": host(st.tecr_traj)[:i],
        "E_traj": host(st.E_traj)[:i],
    }
