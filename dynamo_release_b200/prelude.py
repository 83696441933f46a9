"""SparseVFC prelude (reference scVectorField.py:396-415): valid rows, unique rows, control-point sampling,
bandwidth.  These steps define the control points bit-exactly, so they reproduce the reference's
index semantics:

  * valid_ind  = rows whose velocity is finite                                   (:397-398)
  * np.unique(X, axis=0, return_index=True): rows in lexicographic order, index of the FIRST occurrence
    of each distinct row (:403) -- done here as a device-side stable lexicographic row sort (sort by the
    first column, then refine only the tied groups), the 3.7 s host step of the reference at 1M cells;
  * sample_by_velocity (tools/sampling.py:231-246): numpy Generator.choice(p=|v|/sum|v|, replace=False) --
    the random stream is numpy's own, so this one call stays on the host (it touches M draws and one
    cumulative sum);
  * bandwidth_selector (:121-135): exact k-nearest-neighbour mean distance over the M control points.

torch is used for the sorting / gathering primitives on whatever device X lives on (CUDA in the product).
"""
from __future__ import annotations

import numpy as np
import torch


def valid_rows(Y: torch.Tensor) -> torch.Tensor:
    """Indices of rows with finite velocity (reference: np.where(np.isfinite(Y.sum(1)))[0])."""
    return torch.nonzero(torch.isfinite(Y.sum(1)), as_tuple=False).reshape(-1)


def _lexsort_rows(X: torch.Tensor) -> torch.Tensor:
    """Stable lexicographic argsort of the rows of X (column 0 most significant)."""
    n, d = X.shape
    perm = torch.argsort(X[:, 0], stable=True)
    if d == 1 or n < 2:
        return perm
    k0 = X[perm, 0]
    tied = torch.zeros(n, dtype=torch.bool, device=X.device)
    eq = k0[1:] == k0[:-1]
    tied[1:] |= eq
    tied[:-1] |= eq
    pos = torch.nonzero(tied, as_tuple=False).reshape(-1)
    if pos.numel() == 0:
        return perm
    # refine only the tied positions: LSD stable sorts over columns d-1 .. 1, then by the column-0 group
    sub = perm[pos]
    grp = torch.cumsum(torch.cat([torch.zeros(1, dtype=torch.int64, device=X.device),
                                  (k0[pos][1:] != k0[pos][:-1]).to(torch.int64)]), 0)
    order = torch.arange(sub.numel(), device=X.device)
    # the entries of `sub` are already in stable (original-index) order inside each group
    for c in range(d - 1, 0, -1):
        order = order[torch.argsort(X[sub[order], c], stable=True)]
    order = order[torch.argsort(grp[order], stable=True)]
    perm = perm.clone()
    perm[pos] = sub[order]
    return perm


def unique_rows(X: torch.Tensor):
    """np.unique(X, axis=0, return_index=True): (sorted unique rows, index of first occurrence)."""
    perm = _lexsort_rows(X)
    Xs = X[perm]
    if X.shape[0] == 0:
        return Xs, perm
    new = torch.ones(X.shape[0], dtype=torch.bool, device=X.device)
    new[1:] = (Xs[1:] != Xs[:-1]).any(1)
    # stable sort => the first element of every run is the smallest original index, as np.unique returns
    keep = torch.nonzero(new, as_tuple=False).reshape(-1)
    return Xs[keep], perm[keep]


def velocity_weighted_choice(speed: np.ndarray, M: int, seed: int) -> np.ndarray:
    """tools/sampling.py:242-245: default_rng(seed).choice(arange(n), size=M, p=speed/sum(speed), replace=False).

    The random stream is numpy's own (Generator.random); the rest restates Generator.choice's replace=False
    algorithm (numpy/random/_generator.pyx: draw uniforms, cdf = cumsum(p) / cdf[-1], searchsorted(side='right'),
    keep first occurrences, zero the found entries and repeat) with the O(n) passes fused into one C helper
    (dvf_host_cdf: the same strictly sequential double-precision accumulation as numpy.cumsum), so the chosen
    cells are identical -- tests/test_prelude_cpu.py checks it against Generator.choice itself."""
    from . import _lib
    import ctypes as C
    speed = np.ascontiguousarray(speed, dtype=np.float64)
    n = speed.shape[0]
    rng = np.random.default_rng(seed)
    p = speed / np.sum(speed)
    lib = _lib.load()
    cdf = np.empty(n, dtype=np.float64)
    found = np.zeros(M, dtype=np.int64)
    n_uniq = 0
    while n_uniq < M:
        x = rng.random((M - n_uniq,))
        if n_uniq > 0:
            p[found[0:n_uniq]] = 0
        _lib.check(lib.dvf_host_cdf(p.ctypes.data_as(C.c_void_p), n, cdf.ctypes.data_as(C.c_void_p)))
        new = cdf.searchsorted(x, side="right")
        _, unique_indices = np.unique(new, return_index=True)
        unique_indices.sort()
        new = new.take(unique_indices)
        found[n_uniq:n_uniq + new.size] = new
        n_uniq += new.size
    return found


def bandwidth(ctrl) -> float:
    """scVectorField.py:121-135: h = sqrt(2) * mean(dist to the k nearest neighbours) / 1.5,
    k + 1 = max(2, int(0.2 M)).  Exact brute force over the M x M distances (the reference's kd/ball tree is
    exact too, tools/connectivity.py:565-577).  Squared distances are accumulated dimension by dimension with
    separate multiply / add (the order and rounding sklearn's trees use) on the device `ctrl` lives on; only the
    M x (k+1) sorted neighbour distances return to the host for NumPy's mean."""
    c = torch.as_tensor(ctrl, dtype=torch.float64)
    n = c.shape[0]
    k = max(2, int(0.2 * n)) - 1
    ct = c.T.contiguous()
    d2 = torch.zeros((n, n), dtype=torch.float64, device=c.device)
    diff = torch.empty_like(d2)
    for d in range(ct.shape[0]):
        torch.sub(ct[d][:, None], ct[d][None, :], out=diff)
        torch.mul(diff, diff, out=diff)
        d2 += diff
    kk = min(k + 1, n)
    part = torch.sqrt(torch.topk(d2, kk, dim=1, largest=False, sorted=True).values).cpu().numpy()
    return float(np.sqrt(2) * (np.mean(part[:, 1:]) / 1.5))


def run(X: torch.Tensor, Y: torch.Tensor, M: int, seed: int = 0, beta=None, velocity_based_sampling: bool = True):
    """Returns valid_ind (tensor), Xv, Yv (tensors of the valid rows), ctrl_pts (numpy M x D), idx (numpy), beta."""
    valid = valid_rows(Y)
    all_valid = valid.numel() == Y.shape[0]
    Xv = X if all_valid else X[valid]
    Yv = Y if all_valid else Y[valid]
    uniq, first = unique_rows(Xv)
    n_u = uniq.shape[0]
    M = min(int(M), n_u)
    if velocity_based_sampling:
        speed = torch.linalg.norm(Yv[first], dim=1).cpu().numpy()
        idx = velocity_weighted_choice(speed, M, seed)
    else:
        idx = np.random.RandomState(seed=seed).permutation(n_u)[:M]  # scVectorField.py:409-410
    ctrl_t = uniq[torch.as_tensor(idx, device=X.device)]
    ctrl = ctrl_t.cpu().numpy()
    if beta is None:
        h = bandwidth(ctrl_t)
        beta = 1.0 / h**2
    return valid, Xv, Yv, ctrl, idx, float(beta)
