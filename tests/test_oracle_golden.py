"""CPU: the NumPy restatement (oracle/vfc_oracle.py) against the golden vectors produced by executing the
unmodified reference (oracle/gen_golden.py).  This is what pins the oracle."""
import numpy as np
import pytest

from golden_util import CASES, Golden, rel
from oracle import vfc_oracle as o


@pytest.mark.parametrize("name", CASES)
def test_oracle_fit_matches_reference(name):
    g = Golden(name)
    tr = []
    vf = o.sparse_vfc(g.X, g.V, trace=tr, **g.params)
    s = g.stride
    assert vf["iteration"] == int(g["final_iteration"])
    assert np.array_equal(vf["ctrl_idx"], g["final_ctrl_idx"])
    assert np.array_equal(vf["X_ctrl"], g["final_X_ctrl"])
    assert rel(vf["beta"], g["final_beta"]) < 1e-14
    # same BLAS / LAPACK calls in the same order: identical up to the last bits
    assert rel(vf["C"], g["final_C"]) < 1e-12
    assert rel(vf["P"][::s], g["final_P"]) < 1e-12
    assert rel(vf["sigma2"], g["final_sigma2"]) < 1e-12
    assert rel(vf["E_traj"], g["final_E_traj"]) < 1e-12
    assert rel([t["sigma2_in"] for t in tr], g["trace_sigma2_in"]) < 1e-12
    for it in g.kept_iters():
        assert g.lhs_check(it, tr[it]["lhs"]) < 1e-13
        assert rel(tr[it]["rhs"], g["trace_rhs_%d" % it]) < 1e-13
        assert rel(tr[it]["C"], g["trace_C_%d" % it]) < 1e-12
    full = o.train(g.X, g.V, **g.params)
    assert rel(full["V"][::s], g["final_Vfit"]) < 1e-12


@pytest.mark.parametrize("name", CASES)
def test_oracle_calculus_matches_reference(name):
    g = Golden(name)
    vf = g.vf()
    xs = g.X[g["calc_cells"]]
    nj = g["calc_J"].shape[2]
    assert rel(o.field_eval(xs, vf), g["calc_f"]) < 1e-13
    assert rel(o.jacobian(xs[:nj], vf), g["calc_J"]) < 1e-13
    assert rel(o.jacobian(xs[0], vf), g["calc_J_single"]) < 1e-13
    assert rel(o.divergence(xs, vf), g["calc_div"]) < 1e-13
    _, _, an, a = o.acceleration(xs, vf)
    assert rel(an, g["calc_acc"]) < 1e-12 and rel(a, g["calc_acc_mat"]) < 1e-12
    kn, kv = o.curvature(xs, vf, formula=2)
    assert rel(kn, g["calc_curv"]) < 1e-12 and rel(kv, g["calc_curv_mat"]) < 1e-12
    kn1, _ = o.curvature(xs, vf, formula=1)
    assert rel(kn1, g["calc_curv_f1"]) < 1e-12
    Q = g["gene_Q"]
    ng = Q.shape[0]
    Jg = g["calc_J"][:, :, : g["gene_full"].shape[2]]
    assert np.array_equal(o.gene_jacobian(Jg, Q[: ng // 2], Q[ng // 3:]), g["gene_full"])
    assert rel(o.gene_jacobian_element(Jg, Q[1], Q[5]), g["gene_elem"]) < 1e-13
    assert rel(o.to_gene_space(g["calc_acc_mat"], Q), g["gene_vec"]) < 1e-13


def test_oracle_postprocessing_matches_reference():
    """Per-cell angle / speed / determinant restatements against the executed reference (oracle/gen_golden_post.py)."""
    import os
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "post", "post_cells.npz"))
    for D in (2, 3, 30):
        U, V = z["U_%d" % D], z["V_%d" % D]
        a = o.cell_angles(U, V)
        assert np.array_equal(np.isnan(a), np.isnan(z["angle_%d" % D]))
        assert np.array_equal(a[~np.isnan(a)], z["angle_%d" % D][~np.isnan(a)])
        assert np.array_equal(o.speed(V), z["speed_%d" % D])
    g = Golden("hema_n2000_d30_m100")
    Js = o.jacobian(g.X[z["det_cells"]], g.vf())
    # the Jacobians agree to 1e-13; a determinant moves by n * eps * cond relative to that
    cond = np.array([np.linalg.cond(Js[:, :, i]) for i in range(Js.shape[2])])
    assert np.all(np.abs(o.jacobian_det(Js) - z["det"]) <= 1e-12 * cond * np.abs(z["det"]))


def test_oracle_second_derivatives_match_reference():
    """Hessian / Laplacian / curl / torsion / sensitivity restatements against the executed reference."""
    import os
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "post", "post_hessian.npz"))
    g = Golden("ragged_n1200_d12_m64")
    vf, xs = g.vf(), g.X[z["cells12"]]
    Hs = o.hessian_cells(xs, vf)
    assert rel(Hs, z["Hs12"]) < 1e-13
    assert rel(o.laplacian_of(Hs), z["L12"]) < 1e-13 and rel(o.laplacian_of(Hs[:, :, :, 0]), z["L12_single"]) < 1e-13
    assert rel(o.sensitivity_cells(xs, vf), z["S12"]) < 1e-10
    for D in (2, 3):
        vfd = {"X_ctrl": z["Xc_%d" % D], "C": z["C_%d" % D], "beta": 0.4}
        assert rel(o.curl_cells(z["X_%d" % D], vfd), z["curl_%d" % D]) < 1e-13
    vf3 = {"X_ctrl": z["Xc_3"], "C": z["C_3"], "beta": 0.4}
    assert rel(o.torsion_cells(z["X_3"], vf3), z["torsion_3"]) < 1e-12
