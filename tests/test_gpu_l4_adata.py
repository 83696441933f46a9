"""Row a18: the reference's UNMODIFIED adata-level wrappers on the B200 backend.

dynamo/vectorfield/VectorField.py:21-276 (VectorField: restart loop :455-523, angle loop :255-259, uns['VecFld_pca']
:183-193), vector_calculus.py:199-339 (jacobian), :756-809 (acceleration), :812-868 (curvature), :952-1020
(divergence) are loaded from the executed reference (oracle/ref_loader.py: /root/reference here, the staged
oracle/_ref on the GPU box) and run twice on an AnnData stand-in (tests/adata_standin.py): once as they are (NumPy
backend) and once after dynamo_release_b200.patch.install().  EVERY key either run writes to uns / obs / obsm /
layers is compared: names, shapes, dtypes, values (<= 1e-8 relative where the reference reproduces itself to that
level; the per-test floor is measured by refitting the permuted cells with the reference).
"""
import numpy as np
import pytest

from adata_standin import make_pca_adata, snapshot, AnnDataStandIn
from golden_util import rel
from oracle import ref_loader, vfc_oracle as o

pytestmark = pytest.mark.gpu

if not ref_loader.available():   # pragma: no cover  (oracle/_ref ships with the snapshot)
    pytest.skip("executed reference unavailable (neither /root/reference nor oracle/_ref)", allow_module_level=True)


@pytest.fixture(scope="module")
def refmods():
    ref_loader.install()
    VF = ref_loader.load("dynamo.vectorfield.VectorField")
    vc = ref_loader.load("dynamo.vectorfield.vector_calculus")
    return VF, vc


@pytest.fixture()
def b200_backend(refmods):
    from dynamo_release_b200 import patch

    done = patch.install()
    assert ("dynamo.vectorfield.VectorField", "SvcVectorField") in done
    assert ("dynamo.vectorfield.vector_calculus", "SvcVectorField") in done
    yield patch
    patch.uninstall()


# value tolerance per key suffix (relative to max|ref| unless stated); everything not listed must be EQUAL
EXACT = object()


def _compare(a, b, tol, what, loose=()):
    """a: B200 run, b: reference run (snapshots)."""
    assert sorted(a.keys()) == sorted(b.keys()), (what, sorted(set(a) ^ set(b)))
    report = {}
    for k in sorted(b.keys()):
        va, vb = a[k], b[k]
        if vb is None or isinstance(vb, (str, bool, list)):
            assert va == vb, (what, k, va, vb)
            continue
        if isinstance(vb, (int, np.integer)) and not isinstance(vb, bool):
            assert int(va) == int(vb), (what, k, va, vb)
            continue
        va_, vb_ = np.asarray(va), np.asarray(vb)
        assert va_.shape == vb_.shape, (what, k, va_.shape, vb_.shape)
        assert va_.dtype == vb_.dtype, (what, k, va_.dtype, vb_.dtype)
        if vb_.dtype.kind in "iub" or vb_.dtype == object:
            assert np.array_equal(va_, vb_), (what, k)
            continue
        nan_a, nan_b = np.isnan(va_), np.isnan(vb_)
        assert np.array_equal(nan_a, nan_b), (what, k, "NaN pattern")
        if nan_b.all():
            continue
        t = tol
        for suffix, tt in loose:
            if k.endswith(suffix):
                t = tt
        err = rel(np.where(nan_b, 0.0, va_), np.where(nan_b, 0.0, vb_))
        report[k] = err
        assert err <= t, (what, k, err, t)
    return report


def _run_all(VF, vc, ad, regulators, effectors, **vf_kwargs):
    VF.VectorField(ad, basis="pca", method="SparseVFC", **vf_kwargs)
    vc.jacobian(ad, regulators=regulators, effectors=effectors, basis="pca")
    vc.divergence(ad, basis="pca")
    vc.acceleration(ad, basis="pca")
    vc.curvature(ad, basis="pca")
    return snapshot(ad)


def test_l4_wrappers_pca_every_key(refmods, b200_backend):
    VF, vc = refmods
    X, V = o.hematopoiesis_like(2000, 30, seed=0)        # regime in which the reference reproduces itself (1e-9)
    V[[5, 1200]] = np.nan                                 # valid_ind shorter than n_obs: P / angle bookkeeping
    regs, effs = ["g001", "g005", "g007", "g030"], ["g002", "g003", "g011"]
    b200_backend.uninstall()
    ref = _run_all(VF, vc, make_pca_adata(X, V), regs, effs, M=100)
    # the reference's own floor: permuted cells, same control points
    sc = ref_loader.load("dynamo.vectorfield.scVectorField")
    ok = np.isfinite(V.sum(1))
    perm = np.random.default_rng(1).permutation(int(ok.sum()))
    d2 = sc.SvcVectorField(X[ok][perm], V[ok][perm], None, normalize=False, M=100, MaxIter=30).train()
    inv = np.argsort(perm)
    floor = rel(d2["V"][inv], ref["uns/VecFld_pca/V"][ok])
    tol = max(1e-8, 20 * floor)
    b200_backend.install()
    ours = _run_all(VF, vc, make_pca_adata(X, V), regs, effs, M=100)
    rep = _compare(ours, ref, tol, "pca",
                   loose=(("tecr_traj", 1e-6), ("jacobian_det_pca", 1e-6), ("obs_vf_angle_pca", 1e-7),
                          ("jacobian_gene", 5e-7)))
    print("\n[L4 pca] floor relV %.2g, tol %.2g; worst keys: %s"
          % (floor, tol, sorted(rep.items(), key=lambda kv: -kv[1])[:6]))
    # the layout north_star names
    vf = ours  # noqa
    for key in ("uns/VecFld_pca/X_ctrl", "uns/VecFld_pca/C", "uns/jacobian_pca/jacobian", "uns/jacobian_pca/jacobian_gene",
                "obs/jacobian_det_pca", "obs/divergence_pca", "obs/acceleration_pca", "obs/curvature_pca",
                "obsm/acceleration_pca", "obsm/curvature_pca", "layers/acceleration", "layers/curvature",
                "obsm/velocity_pca_SparseVFC", "obsm/X_pca_SparseVFC", "obs/control_point_pca", "obs/inlier_prob_pca",
                "obs/obs_vf_angle_pca"):
        assert key in ours, key
    assert ours["uns/jacobian_pca/jacobian_gene"].dtype == np.float32
    assert ours["uns/jacobian_pca/jacobian"].shape == (30, 30, 2000)
    assert ours["uns/VecFld_pca/<keys>"] == sorted(
        ["X", "valid_ind", "X_ctrl", "ctrl_idx", "Y", "beta", "V", "C", "P", "VFCIndex", "sigma2", "grid", "grid_V",
         "iteration", "tecr_traj", "E_traj", "normalize", "method", "dims"])


def test_l4_jacobian_sampling_single_pair_and_precomputed_divergence(refmods, b200_backend):
    """jacobian(): cell subsampling (vector_calculus.py:259-263), the 1 x 1 elementwise branch (:299-302),
    store_in_adata=False; divergence() reusing the stored Jacobians (:1000-1008) and computing a cell subset."""
    VF, vc = refmods
    X, V = o.hematopoiesis_like(1500, 30, seed=4)
    out = {}
    for backend in ("ref", "b200"):
        (b200_backend.uninstall if backend == "ref" else b200_backend.install)()
        ad = make_pca_adata(X, V)
        VF.VectorField(ad, basis="pca", M=60, restart_num=0)     # restart_num=0: plain train() (:169-170)
        vc.jacobian(ad, regulators=["g004"], effectors=["g009"], sampling="random", sample_ncells=300)
        # divergence() over the same sampled cells: all 300 come from the stored Jacobians (:1000-1008; with MORE cells
        # than stored Jacobians the reference itself indexes out of bounds at :1010, on either backend)
        vc.divergence(ad, basis="pca", sampling="random", sample_ncells=300)
        ret = vc.jacobian(ad, regulators=["g004", "g006"], cell_idx=np.arange(0, 1500, 7), store_in_adata=False)
        s = snapshot(ad)
        s["ret/jacobian_gene"], s["ret/cell_idx"] = ret["jacobian_gene"], ret["cell_idx"]
        del ad.uns["jacobian_pca"]
        s["div_subset"] = vc.divergence(ad, basis="pca", cell_idx=np.arange(3, 900, 11), store_in_adata=False)
        out[backend] = s
    rep = _compare(out["b200"], out["ref"], 1e-8, "sampling",
                   loose=(("tecr_traj", 1e-6), ("jacobian_det_pca", 1e-6), ("obs_vf_angle_pca", 1e-7),
                          ("jacobian_gene", 5e-7)))
    assert out["b200"]["uns/jacobian_pca/jacobian_gene"].shape == (300,)
    print("\n[L4 sampling] worst keys: %s" % sorted(rep.items(), key=lambda kv: -kv[1])[:4])


def test_l4_two_dimensional_basis_with_grid(refmods, b200_backend):
    """basis='umap' (2-D): Grid / grid_V (VectorField.py:125-127 -> scVectorField.py:505-507), Jacobian without a gene
    transform (vector_calculus.py:307-308), cell_idx forced to all cells (:255-256)."""
    VF, vc = refmods
    rng = np.random.default_rng(3)
    n = 1500
    t = rng.uniform(0, 1, n)
    X = np.stack([np.cos(3 * t) * (1 + t), np.sin(3 * t) * (1 + t)], 1) + 0.05 * rng.normal(size=(n, 2))
    V = np.stack([-X[:, 1], X[:, 0]], 1) * 0.5 + 0.02 * rng.normal(size=(n, 2))
    V[:60] = rng.normal(size=(60, 2)) * 2
    out = {}
    for backend in ("ref", "b200"):
        (b200_backend.uninstall if backend == "ref" else b200_backend.install)()
        ad = AnnDataStandIn(np.zeros((n, 4)), obsm={"X_umap": X.copy(), "velocity_umap": V.copy()})
        VF.VectorField(ad, basis="umap", M=20, grid_num=25)
        vc.jacobian(ad, basis="umap")
        vc.divergence(ad, basis="umap")
        vc.acceleration(ad, basis="umap")
        vc.curvature(ad, basis="umap")
        out[backend] = snapshot(ad)
    assert out["b200"]["uns/VecFld_umap/grid"].shape == (625, 2) and out["b200"]["uns/VecFld_umap/grid_V"].shape == (625, 2)
    rep = _compare(out["b200"], out["ref"], 1e-7, "umap",
                   loose=(("tecr_traj", 1e-5), ("obs_vf_angle_umap", 1e-6), ("jacobian_det_umap", 1e-6)))
    print("\n[L4 umap/grid] worst keys: %s" % sorted(rep.items(), key=lambda kv: -kv[1])[:4])


def test_untested_surface_of_round_1(refmods, b200_backend):
    """con_K(return_d=True) and timeit= (utils.py:179-212), velocity_based_sampling=False (scVectorField.py:409-410),
    SvcVectorField.from_adata / get_data (:577-597), VecFldDict key set (:509-532)."""
    import dynamo_release_b200 as dvf

    VF, vc = refmods
    b200_backend.uninstall()
    sc = ref_loader.load("dynamo.vectorfield.scVectorField")
    vfu = ref_loader.load("dynamo.vectorfield.utils")
    rng = np.random.default_rng(8)
    x, y = rng.normal(size=(57, 7)), rng.normal(size=(13, 7))
    K_ref, D_ref = vfu.con_K(x, y, 0.3, return_d=True)
    K, Dm = dvf.con_K(x, y, 0.3, return_d=True)
    assert K.shape == K_ref.shape and Dm.shape == D_ref.shape == (57, 7, 13)
    assert rel(K, K_ref) < 1e-14 and np.array_equal(Dm, D_ref)
    K1, D1 = dvf.con_K(x[:1], y, 0.3, return_d=True)
    K1r, D1r = vfu.con_K(x[:1], y, 0.3, return_d=True)
    assert K1.shape == K1r.shape and D1.shape == D1r.shape and rel(K1, K1r) < 1e-14
    # @timeit-wrapped names accept and drop timeit= (tools/utils.py:785-796 only logs the time)
    Kt = dvf.con_K(x, y, 0.3, timeit=True)
    assert np.array_equal(Kt, dvf.con_K(x, y, 0.3)) and np.array_equal(dvf.con_K(x, y, 0.3, timeit=False), Kt)
    assert rel(dvf.vector_field_function(x, {"X_ctrl": y, "beta": 0.3, "C": y * 0.5}, timeit=True),
               vfu.vector_field_function(x, {"X_ctrl": y, "beta": 0.3, "C": y * 0.5}, timeit=True)) < 1e-13

    # float32 inputs (utils.py:204-205: the return_d path does not upcast): dtypes as the reference, values to float32
    vf32 = {"X_ctrl": y.astype(np.float32), "C": (y * 0.5).astype(np.float32), "beta": 0.3}
    x32 = x.astype(np.float32)
    for args, kw32 in (((x32[0],), {}), ((x32,), {}), ((x32,), {"vectorize": True}), ((x32.astype(np.float64),), {}),
                       ((x32[0],), {"_beta": np.float64(0.3)})):
        d32 = dict(vf32, beta=kw32.pop("_beta")) if "_beta" in kw32 else vf32
        a, b = dvf.Jacobian_rkhs_gaussian(*args, d32, **kw32), vfu.Jacobian_rkhs_gaussian(*args, d32, **kw32)
        assert a.dtype == b.dtype and a.shape == b.shape, (a.dtype, b.dtype, kw32)
        assert rel(a, b) < 5e-6
    assert dvf.vector_field_function(x32, vf32).dtype == vfu.vector_field_function(x32, vf32).dtype == np.float64

    X, V = o.hematopoiesis_like(1800, 30, seed=2)
    kw = dict(M=80, MaxIter=30, velocity_based_sampling=False, seed=3)
    ref = sc.SparseVFC(X, V, None, **kw)
    got = dvf.SparseVFC(X, V, None, **kw)
    assert set(got.keys()) == set(ref.keys())
    assert np.array_equal(got["ctrl_idx"], ref["ctrl_idx"]) and np.array_equal(got["X_ctrl"], ref["X_ctrl"])
    assert got["iteration"] == ref["iteration"]
    assert rel(got["V"], ref["V"]) < 1e-8 and rel(got["P"], ref["P"]) < 1e-8
    for k in ref:
        assert type(got[k]) is type(ref[k]) or (np.isscalar(got[k]) and np.isscalar(ref[k])), (k, type(got[k]), type(ref[k]))

    # from_adata / get_data on a stored fit
    ad = make_pca_adata(X, V)
    got2 = dict(got, method="SparseVFC", normalize=False)
    ad.uns["VecFld_pca"] = got2
    ours, theirs = dvf.SvcVectorField(), sc.SvcVectorField()
    ours.from_adata(ad, basis="pca")
    theirs.from_adata(ad, basis="pca")
    assert np.array_equal(ours.get_data()[0], theirs.get_data()[0]) and np.array_equal(ours.get_data()[1], theirs.get_data()[1])
    assert ours.vf_dict is ad.uns["VecFld_pca"]
    xs = X[:50]
    assert rel(ours.func(xs), theirs.func(xs)) < 1e-12
    assert rel(ours.get_Jacobian()(xs), theirs.get_Jacobian()(xs)) < 1e-11
    assert np.array_equal(ours.get_X([1, 5]), theirs.get_X([1, 5])) and np.array_equal(ours.get_V(3), theirs.get_V(3))
