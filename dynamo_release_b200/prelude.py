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


_PINNED = {}


def host_f64(t: torch.Tensor) -> np.ndarray:
    """Device vector -> host NumPy array through a cached pinned staging buffer (a pageable .cpu() of 8M speeds
    costs more than the sort that produced them)."""
    if not t.is_cuda:
        return t.numpy()
    n = t.numel()
    buf = _PINNED.get("buf")
    if buf is None or buf.numel() < n:
        buf = _PINNED["buf"] = torch.empty(max(n, 1 << 20), dtype=torch.float64, pin_memory=True)
    buf[:n].copy_(t, non_blocking=True)
    torch.cuda.current_stream().synchronize()
    return buf[:n].numpy()


def valid_rows(Y: torch.Tensor) -> torch.Tensor:
    """Indices of rows with finite velocity (reference: np.where(np.isfinite(Y.sum(1)))[0])."""
    return torch.nonzero(torch.isfinite(Y.sum(1)), as_tuple=False).reshape(-1)


def unique_first(key0: torch.Tensor, rows_of) -> torch.Tensor:
    """Indices of the FIRST occurrence of every distinct row, in lexicographic row order (np.unique(axis=0,
    return_index=True)[1]).  Only column 0 of the rows (`key0`) is sorted; full rows are needed just for the entries
    whose first coordinate is tied with a neighbour, and `rows_of(index tensor) -> rows` is called exactly once for
    those (possibly with no indices) -- locally it is an indexing, sharded over ranks it is a small exchange
    (dist.py), which is why the row matrix itself never has to be gathered."""
    n = key0.numel()
    dev_ = key0.device
    perm = torch.argsort(key0, stable=True)
    k0 = key0[perm]
    tied = torch.zeros(n, dtype=torch.bool, device=dev_)
    if n > 1:
        eq = k0[1:] == k0[:-1]
        tied[1:] |= eq
        tied[:-1] |= eq
    pos = torch.nonzero(tied, as_tuple=False).reshape(-1)
    sub = perm[pos]
    Xt = rows_of(sub)
    keep = torch.ones(n, dtype=torch.bool, device=dev_)
    if pos.numel():
        kt = k0[pos]
        grp = torch.cumsum(torch.cat([torch.zeros(1, dtype=torch.int64, device=dev_),
                                      (kt[1:] != kt[:-1]).to(torch.int64)]), 0)
        # the entries of `sub` are in stable (original-index) order inside each group: LSD stable sorts over the
        # columns d-1 .. 1, then by the column-0 group, leave equal rows in original-index order
        order = torch.arange(sub.numel(), device=dev_)
        for c in range(Xt.shape[1] - 1, 0, -1):
            order = order[torch.argsort(Xt[order, c], stable=True)]
        order = order[torch.argsort(grp[order], stable=True)]
        perm = perm.clone()
        perm[pos] = sub[order]
        Xs, gs = Xt[order], grp[order]
        dup = (Xs[1:] == Xs[:-1]).all(1) & (gs[1:] == gs[:-1])
        keep[pos[1:][dup]] = False
    return perm[torch.nonzero(keep, as_tuple=False).reshape(-1)]


def unique_rows(X: torch.Tensor):
    """np.unique(X, axis=0, return_index=True): (sorted unique rows, index of first occurrence)."""
    if X.shape[0] == 0:
        return X, torch.zeros(0, dtype=torch.int64, device=X.device)
    first = unique_first(X[:, 0].contiguous(), lambda i: X[i])
    return X[first], first


def _pairwise_rows(sq, lo, n):
    """NumPy's pairwise_sum over rows lo .. lo+n-1 of sq (one row per summand).  Module level on purpose: as a nested
    recursive function it formed a reference cycle (function <-> its own closure cell) that kept `sq`, the size of
    the whole velocity matrix, alive until the cyclic collector ran -- the caching allocator then saw a different
    set of free blocks every step and went to cudaMalloc at random moments (tens of ms of idle GPU)."""
    if n < 8:
        res = torch.zeros_like(sq[0])
        for i in range(n):
            res = res + sq[lo + i]
        return res
    if n <= 128:
        r = [sq[lo + j].clone() for j in range(8)]
        i = 8
        while i < n - (n % 8):
            for j in range(8):
                r[j] += sq[lo + i + j]
            i += 8
        res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]))
        while i < n:
            res = res + sq[lo + i]
            i += 1
        return res
    n2 = n // 2
    n2 -= n2 % 8
    return _pairwise_rows(sq, lo, n2) + _pairwise_rows(sq, lo + n2, n - n2)


def row_norms(Y: torch.Tensor) -> torch.Tensor:
    """np.linalg.norm(Y, axis=1) = sqrt(add.reduce(Y*Y, axis=1)) with NumPy's summation order along a contiguous
    row (pairwise_sum: straight loop below 8 terms, else 8 interleaved partial sums combined as
    ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)) plus a sequential tail, halves split recursively above 128 terms), done
    with element-wise column operations so a row's result does not depend on how many rows there are -- a shard
    and the whole matrix give the same bits."""
    Yt = Y.T.contiguous()
    sq = Yt * Yt
    ss = _pairwise_rows(sq, 0, Yt.shape[0])
    # CUDA's double sqrt is correctly rounded like NumPy's; torch's vectorised CPU sqrt is not (CPU = tests only)
    return torch.sqrt(ss) if ss.is_cuda else torch.from_numpy(np.sqrt(ss.numpy()))


_HOST_BUF = {}


def _host_buffer(name: str, n: int) -> np.ndarray:
    """Reusable host scratch (fresh 64 MB arrays cost more in page faults than the pass that fills them)."""
    buf = _HOST_BUF.get(name)
    if buf is None or buf.shape[0] < n:
        buf = _HOST_BUF[name] = np.empty(max(n, 1 << 16), dtype=np.float64)
    return buf[:n]


def numpy_sum(a: np.ndarray) -> float:
    """np.sum of a contiguous float64 vector, bit for bit (NumPy's pairwise tree), on a few threads."""
    from . import _lib
    import ctypes as C
    out = C.c_double(0.0)
    _lib.check(_lib.load().dvf_host_pairwise_sum(a.ctypes.data_as(C.c_void_p), a.shape[0], C.byref(out)))
    return out.value


def velocity_weighted_choice(speed: np.ndarray, M: int, seed: int) -> np.ndarray:
    """tools/sampling.py:242-245: default_rng(seed).choice(arange(n), size=M, p=speed/sum(speed), replace=False).

    The random stream is numpy's own (Generator.random); the rest restates Generator.choice's replace=False
    algorithm (numpy/random/_generator.pyx: draw uniforms, cdf = cumsum(p) / cdf[-1], searchsorted(side='right'),
    keep first occurrences, zero the found entries and repeat) with its O(n) passes reduced to what the result
    needs, each bit-identical to NumPy's: total = np.sum(speed) by the same pairwise tree (dvf_host_pairwise_sum),
    the cumulative sum of speed / total with the found cells skipped, every partial sum rounded as in NumPy's
    sequential recurrence but formed on several threads (dvf_host_choice_cumsum -> csrc/host_cumsum.cpp;
    p itself is never stored), and the bisection evaluated on raw[i] / raw[-1] on the fly
    (dvf_host_searchsorted_div; the normalised array is never formed).  The chosen cells are identical --
    tests/test_prelude_cpu.py checks them against Generator.choice itself."""
    from . import _lib
    import ctypes as C
    speed = np.ascontiguousarray(speed, dtype=np.float64)
    n = speed.shape[0]
    rng = np.random.default_rng(seed)
    lib = _lib.load()
    total = numpy_sum(speed)
    # Generator.choice's own input checks (numpy/random/_generator.pyx): p must be finite / non-negative / sum to 1, and
    # replace=False needs at least `size` non-zero entries -- without them the loop below would never finish
    if not np.isfinite(total) or not total > 0:
        raise ValueError("probabilities contain NaN")          # p = speed / 0 (or a non-finite speed) is NaN
    if int(np.count_nonzero(speed)) < M:
        raise ValueError("Fewer non-zero entries in p than size")
    raw = _host_buffer("raw", n)
    found = np.zeros(M, dtype=np.int64)
    n_uniq = 0
    while n_uniq < M:
        x = rng.random((M - n_uniq,))
        skip = np.sort(found[0:n_uniq])
        _lib.check(lib.dvf_host_choice_cumsum(speed.ctypes.data_as(C.c_void_p), n, total, skip.ctypes.data_as(C.c_void_p),
                                              n_uniq, raw.ctypes.data_as(C.c_void_p)))
        new = np.empty(x.shape[0], dtype=np.int64)
        _lib.check(lib.dvf_host_searchsorted_div(raw.ctypes.data_as(C.c_void_p), n, x.ctypes.data_as(C.c_void_p), x.shape[0],
                                                 new.ctypes.data_as(C.c_void_p)))
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
    if c.is_cuda and n <= 16384:
        # one kernel: distances + sort per control point (kern_post.cu, k_knn_dists)
        from . import _lib
        from .device import stream_ptr
        kk = min(k + 1, n)
        c = c.contiguous()
        out = torch.empty((n, kk), dtype=torch.float64, device=c.device)
        _lib.check(_lib.load().dvf_knn_dists(c.data_ptr(), n, c.shape[1], kk, out.data_ptr(), stream_ptr()))
        part = out.cpu().numpy()
        return float(np.sqrt(2) * (np.mean(part[:, 1:]) / 1.5))
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
    first = unique_first(Xv[:, 0].contiguous(), lambda i: Xv[i])
    n_u = first.numel()
    M = min(int(M), n_u)
    if velocity_based_sampling:
        idx = velocity_weighted_choice(host_f64(row_norms(Yv)[first]), M, seed)
    else:
        idx = np.random.RandomState(seed=seed).permutation(n_u)[:M]  # scVectorField.py:409-410
    ctrl_t = Xv[first[torch.as_tensor(idx, device=X.device)]]
    ctrl = ctrl_t.cpu().numpy()
    if beta is None:
        h = bandwidth(ctrl_t)
        beta = 1.0 / h**2
    return valid, Xv, Yv, ctrl, idx, float(beta)
