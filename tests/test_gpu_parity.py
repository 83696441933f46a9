"""GPU parity tests proper: the sm_100a kernels, called through the C ABI, against the golden vectors produced
by the executed reference and against the NumPy oracle on seeded inputs.

Tolerances (north_star: 1e-8 relative on coefficients, velocities, Jacobians):
  * evaluation with fixed C and step-level ("teacher-forced") quantities: <= 1e-10 of max|ref| (stated per assert);
  * the M x M solve is compared with the exact solution of the reference's own lhs / rhs (long-double iterative
    refinement): ours must be at least as close to it as the reference's normal-equations result;
  * end-to-end fits: <= max(1e-8, 50 x the reference's own permutation floor stored in the golden file) --
    see DESIGN.md "parity floor".
"""
import numpy as np
import pytest
import torch

from golden_util import CASES, Golden, rel
from oracle import vfc_oracle as o

pytestmark = pytest.mark.gpu


def _np(t):
    return t.detach().cpu().numpy()


@pytest.fixture(scope="module")
def pkg():
    import dynamo_release_b200 as p

    return p


@pytest.mark.parametrize("name", CASES)
def test_calculus_fixed_C_matches_reference(pkg, name):
    g = Golden(name)
    vf = g.vf()
    xs = g.X[g["calc_cells"]]
    nj = g["calc_J"].shape[2]
    assert rel(pkg.vector_field_function(xs, vf), g["calc_f"]) < 1e-12
    assert rel(pkg.Jacobian_rkhs_gaussian(xs[:nj], vf), g["calc_J"]) < 1e-11
    J1 = pkg.Jacobian_rkhs_gaussian(xs[0], vf)
    assert J1.shape == g["calc_J_single"].shape and rel(J1, g["calc_J_single"]) < 1e-11
    f_jac, func = pkg.RkhsJacobian(vf), pkg.RkhsField(vf)
    assert rel(pkg.compute_divergence(f_jac, xs), g["calc_div"]) < 1e-11
    acc, acc_mat = pkg.compute_acceleration(func, f_jac, xs)
    assert rel(acc, g["calc_acc"]) < 1e-10 and rel(acc_mat, g["calc_acc_mat"]) < 1e-10
    curv, curv_mat = pkg.compute_curvature(func, f_jac, xs, formula=2)
    assert rel(curv, g["calc_curv"]) < 1e-9 and rel(curv_mat, g["calc_curv_mat"]) < 1e-9
    curv1, none = pkg.compute_curvature(func, f_jac, xs, formula=1)
    assert none is None and rel(curv1, g["calc_curv_f1"]) < 1e-9
    # generic-callable path (Js supplied) agrees with the fused kernel
    Js = pkg.Jacobian_rkhs_gaussian(xs, vf)
    assert rel(pkg.compute_divergence(None, xs, Js=Js), g["calc_div"]) < 1e-11
    acc2, _ = pkg.compute_acceleration(func, None, xs, Js=Js)
    assert rel(acc2, g["calc_acc"]) < 1e-10
    # single point / con_K shapes (utils.py:196-197)
    assert pkg.vector_field_function(xs[0], vf).shape == (xs.shape[1],)
    K = pkg.con_K(xs, vf["X_ctrl"], vf["beta"])
    assert rel(K, o.gauss_kernel(xs, vf["X_ctrl"], vf["beta"])) < 1e-14
    assert pkg.con_K(xs[:1], vf["X_ctrl"], vf["beta"]).ndim == 1


@pytest.mark.parametrize("name", CASES)
def test_gene_space_transforms(pkg, name):
    g = Golden(name)
    Q = g["gene_Q"]
    ng = Q.shape[0]
    Jg = np.ascontiguousarray(g["calc_J"][:, :, : g["gene_full"].shape[2]])
    out = pkg.transform_jacobian(Jg, Q[: ng // 2], Q[ng // 3:])
    ref = g["gene_full"]
    assert out.dtype == np.float32 and out.shape == ref.shape
    ulp = np.abs(out.astype(np.float64) - ref.astype(np.float64)) / np.maximum(np.spacing(np.abs(ref)), 1e-45)
    assert ulp.max() <= 1.0, "float32 result differs by more than 1 ulp"
    # top-k regulator ranking per effector (what rank_jacobian_genes 'full reg' sorts: mean over cells, rank_vf.py:497)
    # from our float32 array equals the ranking from the reference's array
    k = min(5, ref.shape[1])
    for take_abs in (False, True):
        a, b = (np.abs(out), np.abs(ref)) if take_abs else (out, ref)
        ra = np.argsort(-np.mean(a, axis=2), axis=1, kind="stable")[:, :k]
        rb = np.argsort(-np.mean(b, axis=2), axis=1, kind="stable")[:, :k]
        assert np.array_equal(ra, rb)
    assert rel(pkg.elementwise_jacobian_transformation(Jg, Q[1], Q[5]), g["gene_elem"]) < 1e-12
    assert rel(pkg.vector_transformation(g["calc_acc_mat"], Q), g["gene_vec"]) < 1e-13
    # fused |.| group mean == mean of the materialised array
    lab = np.arange(Jg.shape[2]) % 3
    gm = pkg.average_jacobian_by_group(Jg, Q[: ng // 2], Q[ng // 3:], lab, 3, take_abs=True)
    want = np.stack([np.abs(ref[:, :, lab == k]).astype(np.float64).mean(2) for k in range(3)])
    assert rel(gm, want) < 1e-12


def test_gene_transform_tensor_core_path(pkg):
    """Large gene sets take the two-stage DGEMM path (Ge, Gr >= 96): float32 result within 1 ulp of the float64
    einsum, group sums equal to sums of the materialised float32 array, ragged tiles in every dimension."""
    from dynamo_release_b200 import kernels
    rng = np.random.default_rng(21)
    D, n, Ge, Gr = 50, 2 * 173, 200, 131
    J = rng.normal(size=(D, D, n))
    Qe, Qr = rng.normal(size=(Ge, D)) / 7, rng.normal(size=(Gr, D)) / 7
    ref64 = np.einsum("ei,ijc,rj->erc", Qe, J, Qr, optimize=True)
    ref = ref64.astype(np.float32)
    out, _ = kernels.gene_transform(J, Qe, Qr)
    out = _np(out)
    assert out.dtype == np.float32 and out.shape == ref.shape
    ulp = np.abs(out.astype(np.float64) - ref64) / np.maximum(np.spacing(np.abs(ref)).astype(np.float64), 1e-45)
    assert ulp.max() <= 1.0 + 1e-6, "float32 result differs by more than 1 ulp from the float64 value"
    lab = (np.arange(n) * 7) % 5
    full, gs = kernels.gene_transform(J, Qe, Qr, take_abs=True, group=lab, n_groups=5, full=True)
    _, gs_only = kernels.gene_transform(J, Qe, Qr, take_abs=True, group=lab, n_groups=5, full=False)
    full = _np(full)
    assert np.array_equal(full, np.abs(out))
    want = np.stack([full[:, :, lab == k].astype(np.float64).sum(2) for k in range(5)])
    assert rel(_np(gs), want) < 1e-13 and rel(_np(gs_only), want) < 1e-13


def _exact_solve(A, B):
    """long-double iterative refinement of A C = B (reference lhs/rhs) -- the yardstick for the solver."""
    A128, B128 = A.astype(np.longdouble), B.astype(np.longdouble)
    Cx = np.linalg.solve(A, B).astype(np.longdouble)
    for _ in range(6):
        r = (B128 - A128 @ Cx).astype(np.float64)
        Cx = Cx + np.linalg.solve(A, r).astype(np.longdouble)
    return Cx.astype(np.float64)


@pytest.mark.parametrize("name", CASES)
def test_em_steps_teacher_forced(pkg, name):
    from dynamo_release_b200 import kernels as k

    g = Golden(name)
    p = {"a": 5.0, "gamma": 0.9, "lambda_": 3.0, "minP": 1e-5, "theta": 0.75}
    p.update({kk: v for kk, v in g.params.items() if kk in p})
    valid = np.where(np.isfinite(g.V.sum(1)))[0]
    X, Y = g.X[valid], g.V[valid]
    N, D = X.shape
    Xc, beta = g["final_X_ctrl"], float(g["final_beta"])
    M = Xc.shape[0]
    s = g.stride
    UY = k.build_uy(X, Y, Xc, beta)
    U_ref = o.gauss_kernel(X, Xc, beta)
    assert rel(_np(UY[:, :M]), U_ref) < 1e-14
    assert np.array_equal(_np(UY[:, M:M + D]), Y) and float(UY[:, M + D:].abs().sum()) == 0.0
    Kc = o.gauss_kernel(Xc, Xc, beta)
    r2 = np.sum(Y**2, 1)
    for it in (0, 1):
        sigma2, gamma = float(g["trace_sigma2_in"][it]), float(g["trace_gamma_in"][it])
        P, stats = k.estep(r2, D, sigma2, gamma, p["a"])
        assert rel(_np(P)[::s], g["trace_P_%d" % it].reshape(-1)) < 1e-11
        assert rel(_np(stats)[0], g["trace_E_data"][it]) < 1e-11
        G = _np(k.gram(UY, P, p["minP"], D, M))
        lhs = np.tril(G[:M, :M]) + np.tril(G[:M, :M], -1).T + p["lambda_"] * sigma2 * Kc
        rhs = G[M:M + D, :M].T
        assert rel(lhs, g["trace_lhs_%d" % it]) < 1e-12
        assert rel(rhs, g["trace_rhs_%d" % it]) < 1e-12
        # solver: closer to the exact solution of the reference's system than the reference's own result
        C_gpu, info = k.solve(G, Kc, M, D, p["lambda_"] * sigma2)
        assert info == 0
        C_gpu = _np(C_gpu)
        C_exact = _exact_solve(g["trace_lhs_%d" % it], g["trace_rhs_%d" % it])
        err_gpu, err_ref = rel(C_gpu, C_exact), rel(g["trace_C_%d" % it], C_exact)
        assert err_gpu <= max(1e-9, 2 * err_ref), (err_gpu, err_ref)
        # continue from the REFERENCE's C (teacher forcing): V = U C, r2, sum P r2
        C_ref = g["trace_C_%d" % it]
        V, r2_t, st = k.update_v(UY, C_ref, P, p["minP"], D, M)
        V_ref = U_ref.dot(C_ref)
        assert rel(_np(V), V_ref) < 1e-12
        r2 = np.sum((Y - V_ref) ** 2, 1)
        assert rel(_np(r2_t), r2) < 1e-10
        Pc = np.maximum(_np(P), p["minP"])
        assert rel(_np(st)[0], Pc.dot(r2)) < 1e-11
        if it == 0:
            assert rel(Pc.dot(r2) / (Pc.sum() * D), g["trace_sigma2_in"][1]) < 1e-10


@pytest.mark.parametrize("name", CASES)
def test_fit_end_to_end(pkg, name):
    g = Golden(name)
    fc, fv = g.floor()
    tolC, tolV = max(1e-8, 50 * fc), max(1e-8, 50 * fv)
    vf = pkg.SvcVectorField(g.X.copy(), g.V.copy(), None, normalize=False, **g.params).train()
    s = g.stride
    assert np.array_equal(vf["ctrl_idx"], g["final_ctrl_idx"])
    assert np.array_equal(vf["X_ctrl"], g["final_X_ctrl"])
    assert abs(vf["beta"] - float(g["final_beta"])) <= 1e-13 * vf["beta"]
    if not g.lean:
        assert np.array_equal(vf["valid_ind"], g["final_valid_ind"])
    chaotic = fv > 1e-6 or name.startswith("toggle")
    if not chaotic:
        assert vf["iteration"] == int(g["final_iteration"])
        assert rel(vf["E_traj"], g["final_E_traj"]) < tolV
        assert rel(vf["sigma2"], g["final_sigma2"]) < max(tolV, 1e-8)
        assert rel(vf["C"], g["final_C"]) < tolC
        assert rel(vf["P"][::s], g["final_P"]) < max(tolV, 1e-8)
        if not g.lean:
            assert np.array_equal(vf["VFCIndex"], g["final_VFCIndex"])
    assert vf["P"].shape == (len(np.where(np.isfinite(g.V.sum(1)))[0]), 1)
    assert vf["V"].shape == g.X.shape and vf["normalize"] is False
    dv = rel(vf["V"][::s], g["final_Vfit"])
    print("\n[%s] iters %d (ref %d)  relV %.3g (floor %.3g)  relC %.3g (floor %.3g)"
          % (name, vf["iteration"] + 1, int(g["final_iteration"]) + 1, dv, fv, rel(vf["C"], g["final_C"]), fc))
    assert dv < (tolV if not chaotic else max(50 * fv, 1.0))


def test_fit_is_bitwise_reproducible(pkg):
    X, V = o.hematopoiesis_like(6000, 30, seed=21)
    V *= 0.05
    a = pkg.SparseVFC(X, V, None, M=100, MaxIter=10)
    b = pkg.SparseVFC(X, V, None, M=100, MaxIter=10)
    assert np.array_equal(a["C"], b["C"]) and np.array_equal(a["P"], b["P"]) and a["sigma2"] == b["sigma2"]


def test_large_properties_linearity_and_consistency(pkg):
    """Size-independent properties at a larger size: J v from the Jacobian kernel equals the fused acceleration;
    tr J equals the fused divergence; f is linear in C."""
    rng = np.random.default_rng(5)
    n, D, M = 50000, 30, 500
    x = rng.normal(size=(n, D))
    vf = {"X_ctrl": rng.normal(size=(M, D)), "beta": 0.02, "C": rng.normal(size=(M, D))}
    J = pkg.Jacobian_rkhs_gaussian(x, vf, device_out=True)
    v = pkg.vector_field_function(x, vf, device_out=True)
    f_jac, func = pkg.RkhsJacobian(vf), pkg.RkhsField(vf)
    acc, acc_mat = pkg.compute_acceleration(func, f_jac, x)
    assert rel(_np(torch.einsum("ijn,nj->ni", J, v)), acc_mat) < 1e-11
    assert rel(_np(torch.einsum("iin->n", J)), pkg.compute_divergence(f_jac, x)) < 1e-11
    vf2 = dict(vf, C=2.5 * vf["C"])
    assert rel(pkg.vector_field_function(x, vf2), 2.5 * _np(v)) < 1e-14
    # spot check against the oracle on a slice
    sl = slice(1000, 1040)
    assert rel(_np(J[:, :, sl]), o.jacobian(x[sl], vf)) < 1e-11


def test_jacobian_gemm_form_chunks_and_pool_entry(pkg):
    """GEMM-form Jacobian across a K-chunk boundary (131072 cells), ragged last tile, data far from the origin
    (the shift s is the control-point centroid), D = 50; the workspace entry point and the pool entry point agree
    bit for bit and match the oracle's direct formula."""
    from dynamo_release_b200 import kernels
    from dynamo_release_b200.device import dev, empty, lib, ptr, stream_ptr
    from dynamo_release_b200 import _lib

    rng = np.random.default_rng(11)
    n, D, M = 131072 + 1077, 50, 200
    x = rng.normal(size=(n, D)) * 1.5 + 40.0
    vf = {"X_ctrl": x[rng.choice(n, M, replace=False)] + 0.01, "beta": 0.004, "C": rng.normal(size=(M, D))}
    xd, Xc, Cm = dev(x), dev(vf["X_ctrl"]), dev(vf["C"])
    J = kernels.jacobian(xd, Xc, vf["beta"], Cm)
    J2 = empty((D, D, n))
    _lib.check(lib().dvf_jacobian(ptr(xd), n, D, ptr(Xc), M, float(vf["beta"]), ptr(Cm), ptr(J2), stream_ptr()))
    assert torch.equal(J, J2)
    for sl in (slice(0, 24), slice(131072 - 12, 131072 + 12), slice(n - 24, n)):
        assert rel(_np(J[:, :, sl]), o.jacobian(x[sl], vf)) < 1e-11
    assert bool(torch.isfinite(J).all())


def test_edge_cases(pkg):
    rng = np.random.default_rng(1)
    vf = {"X_ctrl": rng.normal(size=(7, 3)), "beta": 0.3, "C": rng.normal(size=(7, 3))}
    assert pkg.vector_field_function(np.zeros((0, 3)), vf).shape == (0, 3)
    assert pkg.Jacobian_rkhs_gaussian(np.zeros((0, 3)), vf).shape == (3, 3, 0)
    x = rng.normal(size=(5, 3))
    assert rel(pkg.vector_field_function(x.astype(np.float32), vf), o.field_eval(x.astype(np.float32).astype(np.float64), vf)) < 1e-13
    assert rel(pkg.vector_field_function(x, vf, X_ctrl_ind=[0, 2]), o.field_eval(x, vf, ctrl_subset=[0, 2])) < 1e-13
    assert rel(pkg.vector_field_function(x, vf, dim=2), o.field_eval(x, vf, dim=2)) < 1e-13
    with pytest.raises(ValueError):
        pkg.vector_field_function(x, {"X_ctrl": vf["X_ctrl"]})
    with pytest.raises(NotImplementedError):
        pkg.SparseVFC(x, x, None, div_cur_free_kernels=True)
    with pytest.raises(NotImplementedError):
        pkg.SvcVectorField(x, x).get_Jacobian("foo")
    with pytest.raises(Exception):
        pkg.SparseVFC(x, np.zeros_like(x) + 1e-9, None, M=3, ecr=2.0)  # loop cannot start: tecr(1) <= ecr


def test_sharded_fit_matches_single_gpu():
    """N-rank NCCL fit == 1-rank fit (runs only where >= 2 GPUs are visible; `gpurun --gpus 2`)."""
    import os
    import subprocess
    import sys

    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29611", os.path.join(root, "tests", "multi_gpu_check.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    print(out.stdout[-2000:], out.stderr[-2000:])
    assert out.returncode == 0 and "MULTI_GPU_CHECK PASS" in out.stdout
