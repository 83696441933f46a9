"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/post/*.npz by EXECUTING THE UNMODIFIED REFERENCE for the
per-cell post-processing and ranking rows.

Run in the build container (needs /root/reference):   python -m oracle.gen_golden_post

  post_cells.npz : angle() per cell (vectorfield/utils.py:1148-1173, called at VectorField.py:255-259),
                   np.linalg.det per cell (vector_calculus.py:329), speed (vector_calculus.py:192)
  post_hessian.npz: Hessian_rkhs_gaussian per cell (utils.py:589-614, loop of vector_calculus.py:413-417), Laplacian
                   (utils.py:665-687), hessian_transformation / elementwise (:617-661), compute_curl (:887-905),
                   compute_torsion (:823-841), compute_sensitivity (:844-870)
  rank_tables.npz: rank_jacobian_genes(..., return_df=True) (rank_vf.py:440-569) on a stand-in adata whose
                   uns['jacobian_pca']['jacobian_gene'] comes from the reference's own subset_jacobian_transformation
                   of the reference's own Jacobian_rkhs_gaussian; tables are stored as index arrays into the gene list
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import ref_loader  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "post")
RANK_CASES = [(mode, ab, norm, grouped, exd)
              for mode in ("full reg", "full eff", "reg", "eff", "int", "switch")
              for ab in (False, True) for norm in (False, True) for grouped in (False, True)
              for exd in ((False, True) if mode in ("reg", "int") else (False,))
              if not (mode == "switch" and ab)]


class _Obs(dict):
    pass


class _Adata:
    def __init__(self, uns, obs):
        self.uns, self.obs = uns, _Obs(obs)


def rank_case_key(mode, ab, norm, grouped, exd):
    return "%s|abs%d|norm%d|grp%d|exd%d" % (mode.replace(" ", "_"), ab, norm, grouped, exd)


def main():
    from golden_util import Golden

    os.makedirs(OUT, exist_ok=True)
    ref_loader.install()
    utils = ref_loader.load("dynamo.vectorfield.utils")
    rank_vf = ref_loader.load("dynamo.vectorfield.rank_vf")
    rng = np.random.default_rng(20240611)

    # ---- per-cell quantities ----
    out = {}
    for D in (2, 3, 30):
        U, V = rng.standard_normal((300, D)), rng.standard_normal((300, D))
        V[:10] = 2.5 * U[:10]
        V[10:20] = -U[10:20]
        U[20] = 0.0
        V[21] = 0.0
        out["U_%d" % D], out["V_%d" % D] = U, V
        out["angle_%d" % D] = np.array([utils.angle(u.astype("float64"), v.astype("float64")) for u, v in zip(U, V)])
        out["speed_%d" % D] = np.array([np.linalg.norm(i) for i in V])
    g = Golden("hema_n2000_d30_m100")
    vf = g.vf()
    xs = g.X[g["calc_cells"]][:96]
    Js = utils.Jacobian_rkhs_gaussian(xs, vf)
    out["det_cells"] = g["calc_cells"][:96]
    out["det"] = np.array([np.linalg.det(Js[:, :, i]) for i in np.arange(Js.shape[2])])
    np.savez_compressed(os.path.join(OUT, "post_cells.npz"), **out)

    # ---- second derivatives, curl / torsion / sensitivity ----
    h = {}
    g12 = Golden("ragged_n1200_d12_m64")
    vf12 = g12.vf()
    xs12 = g12.X[g12["calc_cells"]][:6]
    Hs = np.zeros([xs12.shape[1]] * 3 + [xs12.shape[0]])
    for ind in range(xs12.shape[0]):
        Hs[:, :, :, ind] = utils.Hessian_rkhs_gaussian(xs12[ind], vf12)          # vector_calculus.py:413-417
    h["cells12"] = g12["calc_cells"][:6]
    h["Hs12"] = Hs
    h["L12"] = utils.Laplacian(Hs)
    h["L12_single"] = utils.Laplacian(Hs[:, :, :, 0])
    q = rng.standard_normal((4, xs12.shape[1]))
    h["q"] = q
    h["ht"] = utils.hessian_transformation(Hs[:, :, :, 1], q[:1], q[1:3], q[2:4])
    h["het"] = utils.elementwise_hessian_transformation(Hs[:, :, :, 1], q[0], q[1], q[2])
    f_jac12 = lambda x: utils.Jacobian_rkhs_gaussian(x, vf12)                    # noqa: E731
    h["S12"] = utils.compute_sensitivity(f_jac12, xs12)
    for D in (2, 3):
        vfD = {"X_ctrl": rng.standard_normal((25, D)), "C": 0.3 * rng.standard_normal((25, D)), "beta": 0.4}
        XD = rng.standard_normal((40, D))
        f_jac = lambda x, vfD=vfD: utils.Jacobian_rkhs_gaussian(x, vfD)          # noqa: E731
        f_vf = lambda x, vfD=vfD: utils.vector_field_function(x, vfD)            # noqa: E731
        h["Xc_%d" % D], h["C_%d" % D], h["X_%d" % D] = vfD["X_ctrl"], vfD["C"], XD
        h["curl_%d" % D] = utils.compute_curl(f_jac, XD)
        if D == 3:
            h["torsion_3"] = utils.compute_torsion(f_vf, f_jac, XD)
    np.savez_compressed(os.path.join(OUT, "post_hessian.npz"), **h)

    # ---- ranking tables ----
    G = 14
    genes = np.array(["g%02d" % i for i in range(G)])
    Q = rng.standard_normal((G, xs.shape[1])) / np.sqrt(xs.shape[1])
    Jg = utils.subset_jacobian_transformation(Js, Q, Q)
    labels = np.array(["b", "a", "c"])[rng.integers(0, 3, Js.shape[2])]
    res = {"Q": Q, "genes": genes, "labels": labels, "cells": g["calc_cells"][:96], "Js": Js}
    name_to_i = {n: i for i, n in enumerate(genes)}
    for case in RANK_CASES:
        mode, ab, norm, grouped, exd = case
        ad = _Adata({"jacobian_pca": {"jacobian_gene": Jg.copy(), "cell_idx": np.arange(Js.shape[2]),
                                      "effectors": list(genes), "regulators": list(genes)}}, {"grp": labels})
        tab = rank_vf.rank_jacobian_genes(ad, groups="grp" if grouped else None, jkey="jacobian_pca", abs=ab, mode=mode,
                                          exclude_diagonal=exd, normalize=norm, return_df=True)
        key = rank_case_key(*case)
        if isinstance(tab, dict):     # full reg / full eff: {group: DataFrame with one column per item}
            res[key + "|groups"] = np.array([str(k) for k in tab.keys()])
            for k, df in tab.items():
                res["%s|%s|idx" % (key, k)] = np.array([[name_to_i[x] for x in df[item]] for item in genes])
                res["%s|%s|val" % (key, k)] = np.array([df[item + "_values"].to_numpy(dtype=np.float64) for item in genes])
        else:
            cols = [c for c in tab.columns if not str(c).endswith("_values")]
            res[key + "|groups"] = np.array([str(c) for c in cols])
            for c in cols:
                if mode in ("int", "switch"):
                    pairs = [s.split(" - ") for s in tab[c]]
                    res["%s|%s|idx" % (key, c)] = np.array([[name_to_i[a], name_to_i[b]] for a, b in pairs])
                else:
                    res["%s|%s|idx" % (key, c)] = np.array([name_to_i[x] for x in tab[c]])
                res["%s|%s|val" % (key, c)] = tab[str(c) + "_values"].to_numpy(dtype=np.float64)
    np.savez_compressed(os.path.join(OUT, "rank_tables.npz"), **res)
    print("wrote", OUT, "rank cases:", len(RANK_CASES))


if __name__ == "__main__":
    main()
