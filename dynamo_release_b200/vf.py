"""Host-side mirror of dynamo's vector-field hot-path functions (numpy in -> numpy out, CUDA underneath).

Every public name here has the reference's name, positional order, defaults, return shape / axis order and
exception types; the arithmetic is done by libdvf_b200 (kernels.py).  Functions the
reference decorates with @timeit (tools/utils.py:777-798) accept and drop a `timeit=` keyword.

Pass `device_out=True` to keep results on the GPU as torch tensors (benchmark / pipeline mode).
"""
from __future__ import annotations

import threading

import numpy as np
import torch

from . import kernels
from .device import dev, empty, host


def _out(t, device_out):
    return t if device_out else host(t)


def _vf_parts(vf_dict):
    for key in ("X_ctrl", "beta", "C"):
        if key not in vf_dict:
            raise ValueError("vf_dict lacks the key '%s' (need X_ctrl, beta, C)" % key)
    if vf_dict.get("div_cur_free_kernels", False):
        raise NotImplementedError("div/curl-free kernels are outside the B200 hot path (Gaussian kernel only)")
    return vf_dict["X_ctrl"], float(vf_dict["beta"]), vf_dict["C"]


def _as2d(x):
    if isinstance(x, torch.Tensor):
        return (x[None, :], True) if x.dim() == 1 else (x, False)
    x = np.asarray(x)
    return (x[None, :], True) if x.ndim == 1 else (x, False)


# ------------------------------------------------------------------------------------------------------
def con_K(x, y, beta=0.1, method="cdist", return_d=False, timeit=False, device_out=False):
    """reference vectorfield/utils.py:180-212.  K (n, M), flattened when n == 1 on the values path (:196-197);
    with return_d also D[n, d, m] = x[n, d] - y[m, d]."""
    K = kernels.kernel_matrix(x, y, beta)
    if return_d or method != "cdist":
        xd, yd = dev(x), dev(y)
        D = xd[:, :, None] - yd.t()[None, :, :]
        K = torch.squeeze(K)
        return (_out(K, device_out), _out(D, device_out)) if return_d else _out(K, device_out)
    if K.shape[0] == 1:
        K = K.reshape(-1)
    return _out(K, device_out)


# Field tracing (integrate.py): dynamo wraps the field in plain lambdas, e.g. `lambda x: scale * vector_field_function(
# x=x, vf_dict=VecFld, dim=dims)` (prediction/fate.py:133-135, scVectorField.py:764-766).  While a trace is open on this
# thread every call of vector_field_function notes which field it evaluated and what it returned, so the integrator can
# recognise such a callable as "scale x this RKHS field" and run it on the device.
_TRACE = threading.local()


def _record_trace(x, vf_dict, dim, X_ctrl_ind, result):
    rec = getattr(_TRACE, "rec", None)
    if rec is not None:
        rec.append((np.array(x, dtype=np.float64, copy=True), vf_dict, dim, X_ctrl_ind, result))


def vector_field_function(x, vf_dict, dim=None, kernel="full", X_ctrl_ind=None, timeit=False, device_out=False,
                          **kernel_kwargs):
    """reference vectorfield/utils.py:73-142:  f(x) = K(x, X_ctrl) C."""
    if kernel not in ("full", "df_kernel", "cf_kernel"):
        raise ValueError(f"the kernel can only be one of {'full', 'df_kernel', 'cf_kernel'}!")
    Xc, beta, C = _vf_parts(vf_dict)
    x2, _ = _as2d(x)
    if X_ctrl_ind is not None:  # :128-132
        Cd = dev(C)
        Cm = torch.zeros_like(Cd)
        ind = torch.as_tensor(np.asarray(X_ctrl_ind), device=Cd.device)
        Cm[ind, :] = Cd[ind, :]
        C = Cm
    v = kernels.eval_field(x2, Xc, beta, C)
    if v.shape[0] == 1:
        v = v.reshape(-1)  # con_K flattens a single row (:196-197), so K.dot(C) is 1-d
    if dim is not None:
        v = v[..., :dim] if np.isscalar(dim) else v[..., dim]
    res = _out(v, device_out)
    if getattr(_TRACE, "rec", None) is not None:
        _record_trace(host(x) if isinstance(x, torch.Tensor) else x, vf_dict, dim, X_ctrl_ind, res)
    return res


class RkhsJacobian:
    """Callable returned by SvcVectorField.get_Jacobian('analytical'): x -> (D, D, n).  It carries vf_dict so
    compute_divergence / acceleration / curvature can take the fused no-Jacobian kernel."""

    def __init__(self, vf_dict, **kwargs):
        self.vf_dict = vf_dict
        self.kwargs = kwargs

    def __call__(self, x, **kw):
        return Jacobian_rkhs_gaussian(x, self.vf_dict, **{**self.kwargs, **kw})


def _jacobian_out_dtype(x, vf_dict, single, vectorize):
    """Result dtype of the reference for NumPy inputs: the per-cell loop writes into a float64 array (:368), but the
    1-d path (:364-366) and the vectorised einsum (:374-377) return what `con_K(return_d=True)` produced, which does not
    upcast (:199-205) -- float32 when x, X_ctrl and C all are.  The device computes in float64 either way; the result
    is rounded to float32 there (the reference's float32 ARITHMETIC is not imitated: values agree to float32 rounding)."""
    if not (single or vectorize):
        return np.float64
    parts = [x, vf_dict["X_ctrl"], vf_dict["C"]]
    if isinstance(vf_dict["beta"], np.generic):
        parts.append(vf_dict["beta"])          # a NumPy float64 scalar upcasts (NEP 50), a Python float does not
    if all(isinstance(p, (np.ndarray, np.generic)) for p in parts):
        return np.result_type(*parts, np.float32)
    return np.float64


def Jacobian_rkhs_gaussian(x, vf_dict, vectorize=False, device_out=False):
    """reference vectorfield/utils.py:351-379.  (D, D, n) for 2-d x, (D, D) for 1-d x."""
    Xc, beta, C = _vf_parts(vf_dict)
    x2, single = _as2d(x)
    dt = np.float64 if device_out else _jacobian_out_dtype(x, vf_dict, single, vectorize)
    if not device_out and not single:
        J = kernels.jacobian_to_host(x2, Xc, beta, C)      # slabs stream to pinned host memory behind the GEMM
        return J if dt == np.float64 else J.astype(dt)
    J = kernels.jacobian(x2, Xc, beta, C)
    if single:
        J = J[:, :, 0]
    J = _out(J, device_out)
    return J if (device_out or dt == np.float64) else J.astype(dt)


def Jacobian_rkhs_gaussian_parallel(x, vf_dict, cores=None, device_out=False):
    """reference vectorfield/utils.py:382-396 (ThreadPool over chunks): one kernel launch here."""
    return Jacobian_rkhs_gaussian(x, vf_dict, device_out=device_out)


# ------------------------------------------------------------------------------------------------------
# divergence / acceleration / curvature   (reference vectorfield/utils.py:699-821)
# ------------------------------------------------------------------------------------------------------
def _fused_ok(vf, f_jac, Js):
    """The fused no-Jacobian kernel applies when J is not supplied and both callables are this package's
    RKHS field (same vf_dict)."""
    if Js is not None or not isinstance(f_jac, RkhsJacobian):
        return None
    d = f_jac.vf_dict
    vd = getattr(vf, "vf_dict", None)
    if vf is not None and vd is not None and vd is not d:
        return None
    if vf is not None and vd is None:
        return None
    return d


class RkhsField:
    """Callable x -> f(x) carrying its vf_dict (what SvcVectorField.func is bound to)."""

    def __init__(self, vf_dict):
        self.vf_dict = vf_dict

    def __call__(self, x, **kw):
        return vector_field_function(x, self.vf_dict, **kw)


def _host_pipeline(X, device_out):
    """numpy in -> numpy out with enough cells to overlap the copies with the kernel (kernels.div_acc_curv_host)."""
    return (not device_out) and (not torch.is_tensor(X)) and np.ndim(X) == 2 and np.shape(X)[0] >= 2 * kernels.DAC_STREAM_CHUNK


def _generic_vJ(vf, f_jac, X, Js):
    """Generic callables (reference semantics): evaluate them, continue on the device."""
    v = dev(vf(X)) if vf is not None else None
    J = dev(f_jac(X) if Js is None else Js)
    return v, J


def compute_divergence(f_jac, X, Js=None, vectorize_size=1000, timeit=False, device_out=False):
    """reference :699-725.  div_n = tr J_n."""
    d = _fused_ok(None, f_jac, Js)
    if d is not None:
        Xc, beta, C = _vf_parts(d)
        if _host_pipeline(X, device_out):
            return kernels.div_acc_curv_host(X, Xc, beta, C, want=("div",))["div"]
        return _out(kernels.div_acc_curv(X, Xc, beta, C, want=("div",))["div"], device_out)
    _, J = _generic_vJ(None, f_jac, X, Js)
    return _out(torch.einsum("iin->n", J), device_out)


def compute_acceleration(vf, f_jac, X, Js=None, return_all=False, timeit=False, device_out=False):
    """reference :770-793.  acce (n,), acce_mat (n, D); with return_all also v (n, D) and J (D, D, n)."""
    d = _fused_ok(vf, f_jac, Js)
    if d is not None:
        Xc, beta, C = _vf_parts(d)
        if not return_all and _host_pipeline(X, device_out):
            r = kernels.div_acc_curv_host(X, Xc, beta, C, want=("acc", "acc_norm"))
            return r["acc_norm"], r["acc"]
        r = kernels.div_acc_curv(X, Xc, beta, C, want=("v", "acc", "acc_norm") if return_all else ("acc", "acc_norm"))
        if return_all:
            J = kernels.jacobian(X, Xc, beta, C)
            return _out(r["v"], device_out), _out(J, device_out), _out(r["acc_norm"], device_out), \
                _out(r["acc"], device_out)
        return _out(r["acc_norm"], device_out), _out(r["acc"], device_out)
    v, J = _generic_vJ(vf, f_jac, X, Js)
    acc = torch.einsum("ijn,nj->ni", J, v)
    accn = torch.linalg.norm(acc, dim=1)
    if return_all:
        return _out(v, device_out), _out(J, device_out), _out(accn, device_out), _out(acc, device_out)
    return _out(accn, device_out), _out(acc, device_out)


def compute_curvature(vf, f_jac, X, Js=None, formula=2, timeit=False, device_out=False):
    """reference :797-821 with curvature_method2 (:752-757) / curvature_method1 (:743-749)."""
    d = _fused_ok(vf, f_jac, Js)
    if d is not None:
        Xc, beta, C = _vf_parts(d)
        want = ("curv", "curv_norm") if formula == 2 else ("curv_norm",)
        if _host_pipeline(X, device_out):
            r = kernels.div_acc_curv_host(X, Xc, beta, C, formula=formula, want=want)
            return r["curv_norm"], r.get("curv")
        r = kernels.div_acc_curv(X, Xc, beta, C, formula=formula, want=want)
        return _out(r["curv_norm"], device_out), (None if formula != 2 else _out(r["curv"], device_out))
    v, J = _generic_vJ(vf, f_jac, X, Js)
    a = torch.einsum("ijn,nj->ni", J, v)
    vv = (v * v).sum(1)
    if formula == 1:
        curv = torch.sqrt(vv * (a * a).sum(1)) / torch.sqrt(vv) ** 3
        return _out(curv, device_out), None
    cm = (a * vv[:, None] - v * (v * a).sum(1)[:, None]) / (torch.sqrt(vv) ** 4)[:, None]
    return _out(torch.linalg.norm(cm, dim=1), device_out), _out(cm, device_out)


# ------------------------------------------------------------------------------------------------------
# PCA -> gene space   (reference vectorfield/utils.py:311-326, 430-554)
# ------------------------------------------------------------------------------------------------------
def vector_transformation(V, Q, device_out=False):
    """reference :311-326:  V Q^T, Q is (genes x PCs)."""
    Vd = dev(V)
    single = Vd.dim() == 1
    out = kernels.vector_transform(Vd[None, :] if single else Vd, Q)
    return _out(out[0] if single else out, device_out)


def elementwise_jacobian_transformation(Js, qi, qj, timeit=False, device_out=False):
    """reference :430-453:  q_i J_n q_j for every cell, float64 (n,)."""
    Jd = dev(Js)
    if Jd.dim() == 2:
        Jd = Jd[:, :, None]
    return _out(kernels.gene_transform_elem(Jd.contiguous(), np.asarray(qi).reshape(-1), np.asarray(qj).reshape(-1)),
                device_out)


def transform_jacobian(Js, Qi, Qj, pbar=False, device_out=False):
    """reference :544-554:  (Ge, Gr, n) float32 = Qi J_n Qj^T."""
    Jd = dev(Js)
    if Jd.dim() == 2:
        Jd = Jd[:, :, None]
    Qi, Qj = np.atleast_2d(np.asarray(Qi)), np.atleast_2d(np.asarray(Qj))
    out, _ = kernels.gene_transform(Jd.contiguous(), Qi, Qj)
    return _out(out, device_out)


def subset_jacobian_transformation(Js, Qi, Qj, cores=1, timeit=False, device_out=False):
    """reference :497-541 (`cores` chunks the cells over a ThreadPool there; one launch here)."""
    return transform_jacobian(Js, Qi, Qj, device_out=device_out)


def average_jacobian_by_group(Js, Qi, Qj, group_labels, n_groups, take_abs=False):
    """Fused consumer of K8: per-group mean of (|.|) of the float32 gene-space Jacobian, without materialising
    the (Ge, Gr, n) array -- what rank_jacobian_genes(groups=...) reduces to (rank_vf.py:483-508,
    utils.py:557-582).  Returns float64 (n_groups, Ge, Gr)."""
    Jd = dev(Js)
    lab = np.asarray(group_labels, dtype=np.int32)
    _, gsum = kernels.gene_transform(Jd.contiguous(), np.atleast_2d(Qi), np.atleast_2d(Qj), take_abs=take_abs,
                                     group=lab, n_groups=int(n_groups), full=False)
    cnt = np.bincount(lab, minlength=int(n_groups)).astype(np.float64)
    return host(gsum) / np.maximum(cnt, 1.0)[:, None, None]


# ------------------------------------------------------------------------------------------------------
# per-cell post-processing the reference's adata wrappers do in Python loops
# ------------------------------------------------------------------------------------------------------
def jacobian_det(Js, device_out=False):
    """det of every cell's Jacobian, Js (D, D, n) -> (n,): the list comprehension of np.linalg.det at
    reference vectorfield/vector_calculus.py:329."""
    Jd = dev(Js)
    if Jd.dim() == 2:
        Jd = Jd[:, :, None]
    return _out(kernels.jacobian_det(Jd.contiguous()), device_out)


def jacobian_det_rkhs(x, vf_dict, device_out=False, chunk=131072):
    """det J(x_c) for every row of x straight from the fitted field: J is produced and consumed on the device one
    chunk of cells at a time, so the (D, D, n) array never exists (vector_calculus.py:265-266 + :329)."""
    Xc, beta, C = _vf_parts(vf_dict)
    x2, _ = _as2d(x)
    xd = dev(x2)
    n = xd.shape[0]
    out = empty((n,))
    for lo in range(0, n, int(chunk)):
        hi = min(n, lo + int(chunk))
        out[lo:hi] = kernels.jacobian_det(kernels.jacobian(xd[lo:hi], Xc, beta, C))
    return _out(out, device_out)


def cell_angles(U, V, device_out=False):
    """Signed angle between the observed (U) and fitted (V) velocity of every cell: the per-cell loop over
    `angle` at reference vectorfield/VectorField.py:255-259 (angle: vectorfield/utils.py:1148-1163)."""
    return _out(kernels.cell_angles(np.asarray(U, dtype=np.float64) if not torch.is_tensor(U) else U,
                                    np.asarray(V, dtype=np.float64) if not torch.is_tensor(V) else V), device_out)


def speed(V, device_out=False):
    """|v| per cell: reference vectorfield/vector_calculus.py:192."""
    return _out(kernels.row_norms(V), device_out)


# ------------------------------------------------------------------------------------------------------
# second derivatives and the low-dimensional calculus helpers
# ------------------------------------------------------------------------------------------------------
def Hessian_rkhs_gaussian(x, vf_dict, device_out=False):
    """reference vectorfield/utils.py:589-614: H[i, l, m] = d^2 f_i / dx_l dx_m at the 1-D point x, (D, D, D).
    Extension: a 2-D x (n, D) returns the (D, D, D, n) array dyn.vf.hessian assembles in its cell loop
    (vector_calculus.py:413-417)."""
    Xc, beta, C = _vf_parts(vf_dict)
    x2, single = _as2d(x)
    H = kernels.hessian(x2, Xc, beta, C)
    H = H[0] if single else H.permute(1, 2, 3, 0).contiguous()
    return _out(H, device_out)


def Laplacian(H):
    """reference vectorfield/utils.py:665-687: sum of the diagonal over the FIRST TWO axes of H, per remaining index;
    (D, n) for the 4-D array of dyn.vf.hessian, (D, 1) for one cell's (D, D, D)."""
    if torch.is_tensor(H):
        L = torch.einsum("aak...->k...", H)
        return L if H.dim() == 4 else L[:, None]
    H = np.asarray(H)
    L = np.einsum("aak...->k...", H)
    return L if H.ndim == 4 else L[:, None]


def laplacian_rkhs(x, vf_dict, device_out=False):
    """Laplacian(Hessian(x)) of the reference for every row of x without forming any Hessian: (D, n) like
    Laplacian() returns for the 4-D array ((D, 1) for a 1-D x)."""
    Xc, beta, C = _vf_parts(vf_dict)
    x2, single = _as2d(x)
    return _out(kernels.laplacian(x2, Xc, beta, C).T.contiguous(), device_out)


def hessian_transformation(H, qi, Qj, Qk):
    """reference :617-640: h = Qj (sum_i q_i H_i..) Qk^T for one effector loading row qi."""
    h = np.einsum("ijk, di -> djk", np.asarray(H), np.atleast_2d(qi))
    return np.atleast_2d(Qj) @ h @ np.atleast_2d(Qk).T


def elementwise_hessian_transformation(H, qi, qj, qk):
    """reference :643-661: the scalar q_j^T (sum_i q_i H_i..) q_k."""
    h = np.einsum("ijk, i -> jk", np.asarray(H), np.asarray(qi))
    return np.asarray(qj) @ h @ np.asarray(qk)


def _jac_on_device(f_jac, X):
    """(D, D, n) CUDA tensor: fused path for our own Jacobian callables, any other callable is evaluated as given."""
    if isinstance(f_jac, RkhsJacobian):
        Xc, beta, C = _vf_parts(f_jac.vf_dict)
        return kernels.jacobian(_as2d(X)[0], Xc, beta, C)
    return dev(np.asarray(f_jac(X)))


def compute_curl(f_jac, X, timeit=False, device_out=False):
    """reference :887-905: 2-D: J[1,0] - J[0,1] per cell, (n,); 3-D: the curl vector per cell, stored as the
    reference stores it -- broadcast over the rows of an (n, 3, 3) array."""
    X = np.asarray(X) if not torch.is_tensor(X) else X
    if X.shape[1] > 3:
        raise Exception("curl is only defined in 2/3 dimension.")
    J = _jac_on_device(f_jac, X)
    if X.shape[1] == 2:
        return _out(J[1, 0] - J[0, 1], device_out)
    c = torch.stack([J[2, 1] - J[1, 2], J[0, 2] - J[2, 0], J[1, 0] - J[0, 1]], dim=1)     # (n, 3)
    return _out(c[:, None, :].expand(-1, 3, -1).contiguous(), device_out)


def compute_torsion(vf, f_jac, X, timeit=False, device_out=False):
    """reference :823-841 with torsion_ (:760-766): tau = outer(v, a) (J a) / |outer(v, a)|_F^2 per cell (3-D only),
    a 3-vector the reference broadcasts over the rows of an (n, 3, 3) array."""
    X = np.asarray(X) if not torch.is_tensor(X) else X
    if X.shape[1] != 3:
        raise Exception("torsion is only defined in 3 dimension.")
    J = _jac_on_device(f_jac, X)                                   # (3, 3, n)
    if isinstance(vf, RkhsField):
        Xc, beta, C = _vf_parts(vf.vf_dict)
        v = kernels.eval_field(_as2d(X)[0], Xc, beta, C)
    else:
        v = dev(np.asarray(vf(X)))
    a = torch.einsum("ijn,nj->ni", J, v)                            # acceleration_ : J v
    Ja = torch.einsum("ijn,nj->ni", J, a)
    va = v[:, :, None] * a[:, None, :]                              # outer(v, a)
    tau = torch.einsum("nij,nj->ni", va, Ja) / (va * va).sum((1, 2))[:, None]
    return _out(tau[:, None, :].expand(-1, 3, -1).contiguous(), device_out)


def compute_sensitivity(f_jac, X, timeit=False, device_out=False):
    """reference :844-870: S_c = (I - J_c)^-1 diag(1 / diag((I - J_c)^-1)), (D, D, n).  The batched inverse is the
    library's (cuSOLVER through torch.linalg.inv); the Jacobians come from the fused kernel."""
    X = np.asarray(X) if not torch.is_tensor(X) else X
    J = _jac_on_device(f_jac, X)
    D = J.shape[0]
    A = torch.eye(D, dtype=torch.float64, device=J.device)[None] - J.permute(2, 0, 1)
    s = torch.linalg.inv(A)
    S = s / torch.diagonal(s, dim1=1, dim2=2)[:, None, :]
    return _out(S.permute(1, 2, 0).contiguous(), device_out)
