"""Device-level wrappers: CUDA tensors in, CUDA tensors out, one C-ABI call each (include/dvf_b200.h)."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from .device import dev, empty, lib, ptr, stream_ptr

MAX_D = 64


def _field(x, Xc, Cm=None):
    x, Xc = dev(x), dev(Xc)
    if x.dim() != 2 or Xc.dim() != 2 or x.shape[1] != Xc.shape[1]:
        raise ValueError("x must be (n, D) and X_ctrl (M, D): got %s and %s" % (tuple(x.shape), tuple(Xc.shape)))
    if Cm is not None:
        Cm = dev(Cm)
        if tuple(Cm.shape) != tuple(Xc.shape):
            raise ValueError("C must have the shape of X_ctrl")
    return x, Xc, Cm


def kernel_matrix(x, Xc, beta):
    """K1 -- reference vectorfield/utils.py:180-212 (con_K values)."""
    x, Xc, _ = _field(x, Xc)
    n, D = x.shape
    M = Xc.shape[0]
    K = empty((n, M))
    _lib.check(lib().dvf_kernel_matrix(ptr(x), n, D, ptr(Xc), M, float(beta), ptr(K), M, stream_ptr()))
    return K


def eval_field(x, Xc, beta, Cm):
    """K5 -- reference vectorfield/utils.py:73-142."""
    x, Xc, Cm = _field(x, Xc, Cm)
    n, D = x.shape
    v = empty((n, D))
    _lib.check(lib().dvf_eval(ptr(x), n, D, ptr(Xc), Xc.shape[0], float(beta), ptr(Cm), ptr(v), stream_ptr()))
    return v


_WS = {}


def _workspace(nbytes):
    """Scratch for the launch about to be issued on the current stream, kept per (device, stream) and grown on demand.
    A fresh torch.empty per call looked free but was not: the caching allocator splits the released 0.5 GB block for
    the next smaller request, so every other call went to cudaMalloc (1 ms, at times 60 ms of idle GPU -- measured
    in bench.py's host log).  Calls on one stream are ordered, so they can share the buffer; release_workspaces()
    gives the memory back."""
    key = (torch.cuda.current_device(), stream_ptr())
    ws = _WS.get(key)
    if ws is None or ws.numel() * 8 < nbytes:
        ws = _WS[key] = torch.empty((nbytes + 7) // 8, dtype=torch.float64, device="cuda")
    return ws


def release_workspaces():
    _WS.clear()


def jacobian(x, Xc, beta, Cm, out=None):
    """K6 -- reference vectorfield/utils.py:351-379; returns (D, D, n)."""
    x, Xc, Cm = _field(x, Xc, Cm)
    n, D = x.shape
    J = empty((D, D, n)) if out is None else out
    M = Xc.shape[0]
    nbytes = C.c_size_t(0)
    _lib.check(lib().dvf_jacobian_workspace_bytes(n, D, M, C.byref(nbytes)))
    ws = _workspace(nbytes.value) if nbytes.value else None
    _lib.check(lib().dvf_jacobian_ws(ptr(x), n, D, ptr(Xc), M, float(beta), ptr(Cm), ptr(J), ptr(ws), nbytes.value,
                                     stream_ptr()))
    return J


JAC_STREAM_CHUNK = 131072     # cells per streamed slab (the GEMM kernel's own K chunk)


def jacobian_to_host(x, Xc, beta, Cm):
    """K6 with the result delivered as a NumPy array (D, D, n): cell slabs are computed into two device buffers in
    turn and copied to pinned host memory on a second stream while the next slab is computed, so the PCIe transfer
    of the 8 D^2 n bytes overlaps the GEMM and the full Jacobian never has to exist in HBM."""
    x, Xc, Cm = _field(x, Xc, Cm)
    n, D = x.shape
    Jh = torch.empty((D, D, n), dtype=torch.float64, pin_memory=True)
    if n == 0:
        return Jh.numpy()
    ch = min(JAC_STREAM_CHUNK, n)
    bufs = [torch.empty((D * D * ch,), dtype=torch.float64, device="cuda") for _ in range(2 if n > ch else 1)]
    main, side = torch.cuda.current_stream(), _side_stream()
    side.wait_stream(main)
    copied = [None, None]
    L = lib()
    for k, c0 in enumerate(range(0, n, ch)):
        nc = min(ch, n - c0)
        b = k % len(bufs)
        if copied[b] is not None:
            main.wait_event(copied[b])           # the slab buffer is free again
        slab = bufs[b][:D * D * nc].view(D, D, nc)
        jacobian(x[c0:c0 + nc], Xc, beta, Cm, out=slab)
        done = torch.cuda.Event()
        done.record(main)
        side.wait_event(done)
        _lib.check(L.dvf_memcpy2d_async(Jh.data_ptr() + c0 * 8, n * 8, slab.data_ptr(), nc * 8, nc * 8, D * D, 1,
                                        side.cuda_stream))
        copied[b] = torch.cuda.Event()
        copied[b].record(side)
    side.synchronize()
    return Jh.numpy()


_SIDE = {}


def _side_stream():
    d = torch.cuda.current_device()
    if d not in _SIDE:
        _SIDE[d] = torch.cuda.Stream(device=d)
    return _SIDE[d]


def div_acc_curv(x, Xc, beta, Cm, formula=2, want=("v", "div", "acc", "acc_norm", "curv", "curv_norm"), out=None):
    """K7 -- reference vectorfield/utils.py:699-821; returns a dict of the requested outputs.  `out`: a dict returned
    by an earlier call for the same n, D, `want` -- its tensors are overwritten instead of allocating new ones."""
    x, Xc, Cm = _field(x, Xc, Cm)
    n, D = x.shape
    shapes = {"v": (n, D), "div": (n,), "acc": (n, D), "acc_norm": (n,), "curv": (n, D), "curv_norm": (n,)}
    if out is not None:
        for k in shapes:
            if (k in want) != (out.get(k) is not None) and not (k == "curv" and formula == 1):
                raise ValueError("out does not hold the requested outputs")
            if out.get(k) is not None and (tuple(out[k].shape) != shapes[k] or out[k].dtype != torch.float64
                                           or not out[k].is_cuda or not out[k].is_contiguous()):
                raise ValueError("out[%r] must be a contiguous float64 CUDA tensor of shape %s" % (k, shapes[k]))
        out = dict(out)
    else:
        out = {k: (empty(shapes[k]) if k in want else None) for k in shapes}
    if formula == 1:
        out["curv"] = None
    _lib.check(lib().dvf_div_acc_curv(ptr(x), n, D, ptr(Xc), Xc.shape[0], float(beta), ptr(Cm), int(formula),
                                      ptr(out["v"]), ptr(out["div"]), ptr(out["acc"]), ptr(out["acc_norm"]),
                                      ptr(out["curv"]), ptr(out["curv_norm"]), stream_ptr()))
    return out


DAC_STREAM_CHUNK = 131072     # cells per pipelined slab of the host-to-host path below


def _in_stream():
    d = ("in", torch.cuda.current_device())
    if d not in _SIDE:
        _SIDE[d] = torch.cuda.Stream(device=d[1])
    return _SIDE[d]


def div_acc_curv_host(X, Xc, beta, Cm, formula=2, want=("acc", "acc_norm")):
    """K7 for a host array X with the requested outputs delivered as NumPy arrays: cell slabs go host -> device on one
    stream, through the fused kernel on the compute stream and device -> pinned host on a third, two slabs in flight,
    so for n >> slab the call costs about max(H2D, kernel, D2H) instead of their sum."""
    Xh = np.ascontiguousarray(np.asarray(X), dtype=np.float64)
    n, D = Xh.shape
    Xc, Cm = dev(Xc), dev(Cm)
    if formula == 1:
        want = tuple(k for k in want if k != "curv")
    shapes = {"v": (n, D), "div": (n,), "acc": (n, D), "acc_norm": (n,), "curv": (n, D), "curv_norm": (n,)}
    res = {k: torch.empty(shapes[k], dtype=torch.float64, pin_memory=True) for k in want}
    if n == 0:
        return {k: v.numpy() for k, v in res.items()}
    ch = min(DAC_STREAM_CHUNK, n)
    nbuf = 2 if n > ch else 1
    xin = [torch.empty((ch, D), dtype=torch.float64, device="cuda") for _ in range(nbuf)]
    dout = [{k: torch.empty((ch,) + shapes[k][1:], dtype=torch.float64, device="cuda") for k in want} for _ in range(nbuf)]
    main, s_in, s_out = torch.cuda.current_stream(), _in_stream(), _side_stream()
    # the slab buffers above come from the caching allocator in `main`'s order: work still queued on `main` may be using
    # the recycled blocks, so the side streams start behind it
    s_in.wait_stream(main)
    s_out.wait_stream(main)
    Xt = torch.from_numpy(Xh)
    computed, drained = [None] * nbuf, [None] * nbuf
    L = lib()
    for k, c0 in enumerate(range(0, n, ch)):
        nc = min(ch, n - c0)
        b = k % nbuf
        if computed[b] is not None:
            s_in.wait_event(computed[b])          # the input slab buffer has been consumed
        with torch.cuda.stream(s_in):
            xin[b][:nc].copy_(Xt[c0:c0 + nc], non_blocking=True)
        loaded = torch.cuda.Event()
        loaded.record(s_in)
        main.wait_event(loaded)
        if drained[b] is not None:
            main.wait_event(drained[b])           # the output slab buffers have been copied out
        o = dout[b]
        _lib.check(L.dvf_div_acc_curv(xin[b].data_ptr(), nc, D, ptr(Xc), Xc.shape[0], float(beta), ptr(Cm), int(formula),
                                      ptr(o.get("v")), ptr(o.get("div")), ptr(o.get("acc")), ptr(o.get("acc_norm")),
                                      ptr(o.get("curv")), ptr(o.get("curv_norm")), main.cuda_stream))
        computed[b] = torch.cuda.Event()
        computed[b].record(main)
        s_out.wait_event(computed[b])
        with torch.cuda.stream(s_out):
            for key in want:
                res[key][c0:c0 + nc].copy_(o[key][:nc], non_blocking=True)
        drained[b] = torch.cuda.Event()
        drained[b].record(s_out)
    s_out.synchronize()
    return {k: v.numpy() for k, v in res.items()}


def jacobian_det(J):
    """K9 -- reference vector_calculus.py:329: det of every cell's Jacobian, J (D, D, n) -> (n,)."""
    J = dev(J)
    D, _, n = J.shape
    out = empty((n,))
    _lib.check(lib().dvf_jacobian_det(ptr(J), n, D, ptr(out), stream_ptr()))
    return out


def cell_angles(U, V):
    """K9 -- reference VectorField.py:255-259 / utils.py:1148-1163: signed angle between rows of U and V."""
    U, V = dev(U), dev(V)
    n, D = U.shape
    if V.shape != U.shape:
        raise ValueError("U and V must have the same shape")
    out = empty((n,))
    _lib.check(lib().dvf_cell_angles(ptr(U), ptr(V), n, D, ptr(out), stream_ptr()))
    return out


def row_norms(V):
    """K9 -- reference vector_calculus.py:192 (speed): Euclidean norm of every row."""
    V = dev(V)
    n, D = V.shape
    out = empty((n,))
    _lib.check(lib().dvf_row_norms(ptr(V), n, D, ptr(out), stream_ptr()))
    return out


def hessian(x, Xc, beta, Cm):
    """K10 -- reference vectorfield/utils.py:589-614 for every row of x: (n, D, D, D), H[c, i, l, m]."""
    x, Xc, Cm = _field(x, Xc, Cm)
    n, D = x.shape
    out = empty((n, D, D, D))
    _lib.check(lib().dvf_hessian(ptr(x), n, D, ptr(Xc), Xc.shape[0], float(beta), ptr(Cm), ptr(out), stream_ptr()))
    return out


def laplacian(x, Xc, beta, Cm):
    """K10 -- sum_a H[c, a, a, k] (reference utils.py:665-687 on the array of vector_calculus.py:413): (n, D)."""
    x, Xc, Cm = _field(x, Xc, Cm)
    n, D = x.shape
    out = empty((n, D))
    _lib.check(lib().dvf_laplacian(ptr(x), n, D, ptr(Xc), Xc.shape[0], float(beta), ptr(Cm), ptr(out), stream_ptr()))
    return out


def gene_transform(J, Qe, Qr, take_abs=False, group=None, n_groups=0, full=True):
    """K8 -- reference vectorfield/utils.py:544-554.  J (D, D, n) -> float32 (Ge, Gr, n) and/or group sums."""
    J, Qe, Qr = dev(J), dev(Qe), dev(Qr)
    D, _, n = J.shape
    Ge, Gr = Qe.shape[0], Qr.shape[0]
    out = empty((Ge, Gr, n), torch.float32) if full else None
    gsum = None
    if group is not None:
        group = dev(group, torch.int32)
        gsum = torch.zeros((n_groups, Ge, Gr), dtype=torch.float64, device="cuda")
    _lib.check(lib().dvf_gene_transform(ptr(J), n, D, ptr(Qe), Ge, ptr(Qr), Gr, int(bool(take_abs)), ptr(group),
                                        int(n_groups), ptr(out), ptr(gsum), stream_ptr()))
    return out, gsum


def gene_transform_reduce(J, Qe, Qr, op="sum", take_abs=False, jmax=None, group=None, n_groups=1):
    """Reductions of the gene-space Jacobian over cells without materialising it (reference rank_vf.py:483-508).
    op: "sum" (per-group sums), "switch" (per-group sums of J * J^T where both are negative), "max" (max |J|).
    Returns float64 (n_groups, Ge, Gr) for the sums, (Ge, Gr) for "max"."""
    import ctypes as C
    J, Qe, Qr = dev(J), dev(Qe), dev(Qr)
    D, _, n = J.shape
    Ge, Gr = Qe.shape[0], Qr.shape[0]
    code = {"sum": 0, "switch": 1, "max": 2}[op]
    if group is None:
        group = torch.zeros(n, dtype=torch.int32, device="cuda")
    group = dev(group, torch.int32)
    if n % 2 or J.data_ptr() % 16:
        # the TMA row pitch (n doubles) must be a multiple of 16 bytes: pad with one cell that belongs to no group
        Jp = torch.zeros((D, D, n + n % 2), dtype=torch.float64, device="cuda")
        Jp[:, :, :n] = J
        group = torch.cat([group, torch.full((n % 2,), -1, dtype=torch.int32, device="cuda")])
        J, n = Jp, n + n % 2
    out = torch.zeros((Ge, Gr) if code == 2 else (int(n_groups), Ge, Gr), dtype=torch.float64, device="cuda")
    if n == 0:
        return out
    jm = None if jmax is None else dev(jmax, torch.float32)
    nbytes = C.c_size_t(0)
    _lib.check(lib().dvf_gene_gemm_workspace_bytes(n, D, Ge, Gr, 0, 1, C.byref(nbytes)))
    ws = torch.empty((nbytes.value + 1023) // 8 + 128, dtype=torch.float64, device="cuda")
    off = (-ws.data_ptr()) % 1024
    _lib.check(lib().dvf_gene_transform_reduce(ptr(J), n, D, ptr(Qe), Ge, ptr(Qr), Gr, int(bool(take_abs)), code, ptr(jm),
                                               ptr(group), int(n_groups), ptr(out), ws.data_ptr() + off, nbytes.value,
                                               stream_ptr()))
    return out


def gene_transform_elem(J, qe, qr):
    """reference vectorfield/utils.py:430-453."""
    J, qe, qr = dev(J), dev(qe), dev(qr)
    D, _, n = J.shape
    out = empty((n,))
    _lib.check(lib().dvf_gene_transform_elem(ptr(J), n, D, ptr(qe), ptr(qr), ptr(out), stream_ptr()))
    return out


def vector_transform(V, Q):
    """reference vectorfield/utils.py:311-326:  V Q^T."""
    V, Q = dev(V), dev(Q)
    n, D = V.shape
    G = Q.shape[0]
    out = empty((n, G))
    _lib.check(lib().dvf_vector_transform(ptr(V), n, D, ptr(Q), G, ptr(out), stream_ptr()))
    return out


# ---- EM pieces ------------------------------------------------------------------------------------
def row_width(M, D):
    return int(_lib.load().dvf_fit_row_width(int(M), int(D)))


def build_uy(X, Y, Xc, beta):
    X, Xc, _ = _field(X, Xc)
    Y = dev(Y)
    N, D = X.shape
    M = Xc.shape[0]
    UY = empty((N, row_width(M, D)))
    _lib.check(lib().dvf_build_uy(ptr(X), ptr(Y), N, D, ptr(Xc), M, float(beta), ptr(UY), stream_ptr()))
    return UY


def estep(r2, D, sigma2, gamma, a):
    """scVectorField.py:197-227.  Returns P (unclamped) and [E, sum P, min nonzero t1, #zeros]."""
    r2 = dev(r2)
    N = r2.shape[0]
    P = empty((N + 2,))
    stats = empty((4,))
    _lib.check(lib().dvf_estep(ptr(r2), N, int(D), float(sigma2), float(gamma), float(a), ptr(P), ptr(stats),
                               stream_ptr()))
    return P[:N], stats


def gram(UY, P, minP, D, M):
    """scVectorField.py:468-476.  Returns G (ldg x ldg): the lower triangle of columns < M is valid (U^T P U and
    Y^T P U); the Y^T P Y corner is not computed."""
    UY = dev(UY)
    N = UY.shape[0]
    Pp = torch.zeros((N + 2,), dtype=torch.float64, device="cuda")
    Pp[:N] = dev(P).reshape(-1)
    ldg = int(lib().dvf_gram_ld(int(D), int(M)))
    G = torch.zeros((ldg, ldg), dtype=torch.float64, device="cuda")
    nbytes = C.c_size_t(0)
    _lib.check(lib().dvf_gram_scratch_bytes(N, int(D), int(M), C.byref(nbytes)))
    scratch = torch.empty((nbytes.value + 7) // 8, dtype=torch.float64, device="cuda")
    _lib.check(lib().dvf_gram(ptr(UY), ptr(Pp), float(minP), N, int(D), int(M), ptr(G), ptr(scratch),
                              nbytes.value, stream_ptr()))
    return G


def solve(G, Kc, M, D, lam_sigma2):
    """(G[:M,:M] + lam_sigma2 Kc) C = G[M:M+D,:M]^T.   Returns C (M x D), info."""
    G, Kc = dev(G), dev(Kc)
    L = empty((M + D, M))
    Cm = empty((M, D))
    info = torch.zeros((1,), dtype=torch.int32, device="cuda")
    _lib.check(lib().dvf_solve(ptr(G), G.shape[1], ptr(Kc), int(M), int(D), float(lam_sigma2), ptr(L), ptr(Cm),
                               ptr(info), stream_ptr()))
    return Cm, int(info.item())


def solve_normal(G, Kc, M, D, lam_sigma2):
    """The reference's solver form, tools/utils.py:2975-2976: C = solve(lhs^T lhs, lhs^T rhs) by LU with partial
    pivoting (lhs = G[:M,:M] + lam_sigma2 Kc, rhs = G[M:M+D,:M]^T).  Returns C (M x D), info."""
    G, Kc = dev(G), dev(Kc)
    nbytes = C.c_size_t(0)
    _lib.check(lib().dvf_solve_normal_workspace_bytes(int(M), int(D), C.byref(nbytes)))
    ws = torch.empty((nbytes.value + 7) // 8, dtype=torch.float64, device="cuda")
    Cm = empty((M, D))
    info = torch.zeros((1,), dtype=torch.int32, device="cuda")
    _lib.check(lib().dvf_solve_normal(ptr(G), G.shape[1], ptr(Kc), int(M), int(D), float(lam_sigma2), ptr(ws),
                                      nbytes.value, ptr(Cm), ptr(info), stream_ptr()))
    return Cm, int(info.item())


def update_v(UY, Cm, P, minP, D, M):
    """scVectorField.py:484-486.  Returns V, r2, sum max(P,minP) r2."""
    UY, Cm, P = dev(UY), dev(Cm), dev(P).reshape(-1)
    N = UY.shape[0]
    V, r2, stats = empty((N, D)), empty((N,)), empty((1,))
    _lib.check(lib().dvf_update_v(ptr(UY), ptr(Cm), ptr(P), float(minP), N, int(D), int(M), ptr(V), ptr(r2),
                                  ptr(stats), stream_ptr()))
    return V, r2, stats
