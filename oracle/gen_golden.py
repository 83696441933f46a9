"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.npz by EXECUTING THE UNMODIFIED REFERENCE.

Run in the build container (needs /root/reference):   python -m oracle.gen_golden
The vectors travel with the repo; nothing on the GPU box reads /root/reference.

Every file records the software tuple the reference ran on (numpy / scipy / sklearn versions,
reference commit) because the reference pins none of them (requirements.txt lower bounds only).

Per case the file holds
  inputs        X, V (float64), fit parameters
  final         the VecFldDict produced by SvcVectorField(...).train()   (scVectorField.py:997-1046)
  trace         per-EM-iteration sigma2/gamma/E/tecr for all iterations and lhs/rhs/C for a few
                (captured by wrapping get_P / lstsq_solver inside the loaded reference module)
  calculus      Jacobian_rkhs_gaussian, compute_divergence/acceleration/curvature on a cell subset
  gene          transform_jacobian / elementwise_jacobian_transformation on a small loading matrix
  floor         the same fit with the cells permuted: what the reference guarantees of itself
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle import ref_loader, vfc_oracle  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def _versions():
    import scipy
    import sklearn

    return "reference@69ba903(v1.5.3) numpy=%s scipy=%s sklearn=%s" % (np.__version__, scipy.__version__,
                                                                      sklearn.__version__)


def run_reference_fit(sc, X, V, keep_iters=(0, 1), **params):
    """Fit with the reference and capture the per-iteration quantities."""
    trace = {"sigma2_in": [], "gamma_in": [], "E_data": [], "lhs": {}, "rhs": {}, "C": {}, "P": {}}
    orig_get_P, orig_solver = sc.get_P, sc.lstsq_solver
    it = {"n": 0}

    def spy_get_P(Y, Vc, sigma2, gamma, a, *rest):
        P, E = orig_get_P(Y, Vc, sigma2, gamma, a, *rest)
        trace["sigma2_in"].append(sigma2)
        trace["gamma_in"].append(gamma)
        trace["E_data"].append(E)
        if it["n"] in keep_iters:
            trace["P"][it["n"]] = P.copy()
        return P, E

    def spy_solver(lhs, rhs, **kw):
        C = orig_solver(lhs, rhs, **kw)
        if it["n"] in keep_iters:
            trace["lhs"][it["n"]] = lhs.copy()
            trace["rhs"][it["n"]] = rhs.copy()
            trace["C"][it["n"]] = C.copy()
        it["n"] += 1
        return C

    sc.get_P, sc.lstsq_solver = spy_get_P, spy_solver
    try:
        vf = sc.SvcVectorField(X.copy(), V.copy(), None, normalize=False, **params)
        d = vf.train()
    finally:
        sc.get_P, sc.lstsq_solver = orig_get_P, orig_solver
    return vf, d, trace


def rel(a, b):
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


LHS_ROWS = 24    # big-M cases keep this many rows of lhs per iteration (plus its column sums) instead of M x M


def make_case(name, X, V, params, n_jac=48, n_calc=200, n_genes=24, gene_cells=32, lean=None, keep_iters=(0, 1),
              lhs_full=True):
    """lean=(n, d, seed, vscale[, manifold_seed]): inputs are NOT stored; tests regenerate them (4 values:
    vfc_oracle.hematopoiesis_like(n, d, seed) * vscale; 5 values: the bench's generator
    dynamo_release_b200.datasets.hematopoiesis_like with that manifold seed) and check the stored checksums.
    lhs_full=False: per kept iteration only LHS_ROWS evenly spaced rows of lhs and its column sums are stored."""
    sc, vfu, _ = ref_loader.hot_path()
    vf, d, tr = run_reference_fit(sc, X, V, keep_iters=keep_iters, **params)
    out = {"meta": np.array(_versions()), "params_keys": np.array(list(params.keys())),
           "params_vals": np.array([float(v) for v in params.values()])}
    if lean is None:
        out["X"], out["V"] = X, V
    else:
        out["gen"] = np.array(lean, dtype=np.float64)
        out["checksum"] = np.array([X.sum(), V.sum(), np.abs(X).sum(), np.abs(V).sum()])
    for k in ("valid_ind", "X_ctrl", "ctrl_idx", "beta", "V", "C", "P", "VFCIndex", "sigma2", "iteration",
              "tecr_traj", "E_traj"):
        val = np.asarray(d[k])
        if lean is not None and k in ("V", "P", "valid_ind", "VFCIndex"):
            if k == "VFCIndex":
                out["final_n_inlier"] = np.array(len(val))
                continue
            if k == "valid_ind":
                out["final_n_valid"] = np.array(len(val))
                continue
            val = val[::37]   # every 37th row only
        out["final_" + ("Vfit" if k == "V" else k)] = val
    out["trace_sigma2_in"] = np.array(tr["sigma2_in"])
    out["trace_gamma_in"] = np.array(tr["gamma_in"])
    out["trace_E_data"] = np.array(tr["E_data"])
    out["trace_kept_iters"] = np.array(sorted(tr["lhs"].keys()))
    for it_no in tr["lhs"]:
        if lhs_full:
            out["trace_lhs_%d" % it_no] = tr["lhs"][it_no]
        else:
            lhs = tr["lhs"][it_no]
            rows = np.linspace(0, lhs.shape[0] - 1, LHS_ROWS).astype(np.int64)
            out["trace_lhs_rows_idx"] = rows
            out["trace_lhs_rows_%d" % it_no] = lhs[rows]
            out["trace_lhs_colsum_%d" % it_no] = lhs.sum(0)
            out["trace_lhs_cond_%d" % it_no] = np.array(np.linalg.cond(lhs))
        out["trace_rhs_%d" % it_no] = tr["rhs"][it_no]
        out["trace_C_%d" % it_no] = tr["C"][it_no]
        out["trace_P_%d" % it_no] = tr["P"][it_no] if lean is None else tr["P"][it_no][::37]

    rng = np.random.default_rng(1234)
    cells = np.sort(rng.choice(len(X), size=min(n_calc, len(X)), replace=False))
    xs = X[cells]
    out["calc_cells"] = cells
    J = vfu.Jacobian_rkhs_gaussian(xs[:n_jac], d)
    out["calc_J"] = J
    out["calc_J_single"] = vfu.Jacobian_rkhs_gaussian(xs[0], d)
    out["calc_f"] = vfu.vector_field_function(xs, d)
    f_jac = vf.get_Jacobian("analytical")
    out["calc_div"] = vfu.compute_divergence(f_jac, xs)
    acc, acc_mat = vfu.compute_acceleration(vf.func, f_jac, xs)
    out["calc_acc"], out["calc_acc_mat"] = acc, acc_mat
    curv, curv_mat = vfu.compute_curvature(vf.func, f_jac, xs, formula=2)
    out["calc_curv"], out["calc_curv_mat"] = curv, curv_mat
    curv1, _ = vfu.compute_curvature(vf.func, f_jac, xs, formula=1)
    out["calc_curv_f1"] = curv1

    D = X.shape[1]
    Q, _ = np.linalg.qr(np.random.default_rng(77).normal(size=(max(n_genes, D), D)))
    Q = np.ascontiguousarray(Q[:n_genes])
    out["gene_Q"] = Q
    Jg = J[:, :, :gene_cells]
    out["gene_full"] = vfu.transform_jacobian(Jg, Q[: n_genes // 2], Q[n_genes // 3:])
    out["gene_elem"] = vfu.elementwise_jacobian_transformation(Jg, Q[1], Q[5])
    out["gene_vec"] = vfu.vector_transformation(acc_mat, Q)

    # reference-vs-itself floor: permute the cells and refit
    perm = np.random.default_rng(99).permutation(len(X))
    _, d2, tr2 = run_reference_fit(sc, X[perm], V[perm], keep_iters=(), **params)
    out["floor_E_traj"] = np.asarray(d2["E_traj"])
    out["floor_sigma2"] = np.array(d2["sigma2"])
    out["floor_trace_sigma2_in"] = np.array(tr2["sigma2_in"])
    same_ctrl = d2["X_ctrl"].shape == d["X_ctrl"].shape and np.array_equal(d2["X_ctrl"], d["X_ctrl"])
    out["floor_same_ctrl"] = np.array(same_ctrl)
    out["floor_iteration"] = np.array(d2["iteration"])
    if same_ctrl:
        out["floor_relC"] = np.array(rel(d2["C"], d["C"]))
        inv = np.argsort(perm)
        out["floor_relV"] = np.array(rel(d2["V"][inv], d["V"]))
    else:
        out["floor_relC"] = np.array(np.nan)
        out["floor_relV"] = np.array(np.nan)

    # cross-check the NumPy restatement right here, so a drift is caught at generation time
    o = vfc_oracle.train(X, V, **params)
    print("[%s] N=%d D=%d M=%d iters=%d  sigma2=%.6g  floor relC=%.3g relV=%.3g | oracle-vs-ref relC=%.3g relV=%.3g"
          % (name, X.shape[0], X.shape[1], d["C"].shape[0], d["iteration"] + 1, d["sigma2"],
             float(out["floor_relC"]), float(out["floor_relV"]), rel(o["C"], d["C"]), rel(o["V"], d["V"])))
    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)


def toggle_data():
    """Config 1: the reference's own simulator (simulation/ODE.py:523-552 state_space_sampler + :78 toggle)."""
    ode = ref_loader.load("dynamo.simulation.ODE")
    X, Y = ode.state_space_sampler(ode=ode.toggle, dim=2, seed_num=19491001, clip=True, min_val=0, max_val=6, N=5000)
    X, Y = np.asarray(X, dtype=np.float64), np.asarray(Y, dtype=np.float64)
    keep = X.sum(1) > 0  # simulation/ODE.py:641-645
    return np.ascontiguousarray(X[keep]), np.ascontiguousarray((Y - X)[keep])  # velocity layer = Y - X (:638)


def lean_case(name, n, d, seed, vscale, params, **kw):
    X, V = vfc_oracle.hematopoiesis_like(n, d, seed=seed)
    V = V * vscale
    make_case(name, X, V, params, lean=(n, d, seed, vscale), **kw)


def bench_case(name, n, d, seed, vscale, manifold_seed, params, **kw):
    """Cases at the configurations bench.py quotes its numbers on, from bench.py's own generator."""
    from dynamo_release_b200.datasets import hematopoiesis_like

    X, V = hematopoiesis_like(n, d, seed=seed, velocity_scale=vscale, manifold_seed=manifold_seed)
    make_case(name, X, V, params, lean=(n, d, seed, vscale, manifold_seed), keep_iters=(0, 1, 2, 3), lhs_full=False,
              n_jac=24, n_calc=100, gene_cells=16, **kw)


def main():
    only = sys.argv[1:]
    if "bench" in only or not only:
        # BASELINE.json configs[1] exactly (C2), a C3-shaped case (M = 500) and a C5-shaped case (D = 50, M = 1000):
        # bench.py's generator, seed (100 = rank 0) and velocity_scale
        bench_case("c2_n100000_d30_m100", 100000, 30, 100, 0.05, 7, dict(M=100, MaxIter=30))
        bench_case("c3s_n50000_d30_m500", 50000, 30, 100, 0.05, 7, dict(M=500, MaxIter=30))
        bench_case("c5s_n20000_d50_m1000", 20000, 50, 100, 0.05, 7, dict(M=1000, MaxIter=10))
        if only:
            return
    if "lean" in only or not only:
        # non-degenerate EM (sigma2 small enough that P is not clamped everywhere; far outliers hit the t1 == 0 rule)
        lean_case("lean_n20000_d30_m100", 20000, 30, 11, 0.05, dict(M=100, MaxIter=30))
        lean_case("lean_n8000_d50_m200", 8000, 50, 12, 0.05, dict(M=200, MaxIter=30), n_jac=24, n_calc=100,
                  gene_cells=16)
        if only:
            return
    X, V = vfc_oracle.hematopoiesis_like(2000, 30, seed=0)
    make_case("hema_n2000_d30_m100", X, V, dict(M=100, MaxIter=30))
    X, V = vfc_oracle.hematopoiesis_like(1500, 50, seed=3)
    make_case("hema_n1500_d50_m160", X, V, dict(M=160, MaxIter=30), n_jac=24, n_calc=100, gene_cells=16)
    X, V = toggle_data()
    make_case("toggle_n5000_d2_m100", X, V, dict(M=100, MaxIter=30))
    # a case with non-finite velocity rows and duplicated cells (prelude edge cases, scVectorField.py:397-403)
    X, V = vfc_oracle.hematopoiesis_like(1200, 12, seed=5)
    V[[3, 77, 500]] = np.nan
    V[901, 4] = np.inf
    X[100:110] = X[90:100]
    make_case("ragged_n1200_d12_m64", X, V, dict(M=64, MaxIter=12, lambda_=2.0, gamma=0.8), n_jac=16, n_calc=64,
              gene_cells=16, n_genes=12)


if __name__ == "__main__":
    main()
