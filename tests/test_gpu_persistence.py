"""Row f4 (second half): the reference's only "checkpoint" is the saved AnnData.  A fit made by the B200 backend goes
through dynamo's own `cleanup` (data_io.py:287-336), to a file and back, and `vecfld_from_adata`
(vectorfield/utils.py:269-308) / `SvcVectorField.from_adata` rebuild a field from the reloaded dict that evaluates bit
for bit like the one in memory.

h5ad itself is unavailable in this image (no h5py / anndata): the file is an .npz written with h5ad's value model --
one dataset per leaf of the nested `uns` dict, scalars as 0-d datasets that come back as NumPy scalars, strings as str,
`None` leaves not stored at all (what anndata does with them), lists of str as string arrays."""
import os

import numpy as np
import pytest

from adata_standin import make_pca_adata
from oracle import ref_loader, vfc_oracle as o

pytestmark = pytest.mark.gpu


def save_uns(path, uns):
    flat = {}

    def walk(prefix, v):
        if isinstance(v, dict):
            for k, x in v.items():
                walk(prefix + "/" + str(k), x)
        elif v is None:
            return                                      # h5ad has no dataset for None
        else:
            flat[prefix] = np.asarray(v)
    walk("uns", uns)
    np.savez(path, **flat)


def load_uns(path):
    out = {}
    with np.load(path, allow_pickle=False) as z:
        for key in z.files:
            parts = key.split("/")[1:]
            d = out
            for p in parts[:-1]:
                d = d.setdefault(p, {})
            v = z[key]
            if v.ndim == 0:
                v = v[()]                               # 0-d dataset -> NumPy scalar (np.float64, np.int64, np.bool_, np.str_)
                if isinstance(v, np.str_):
                    v = str(v)
            d[parts[-1]] = v
    return out


def test_saved_fit_reloads_into_the_gpu_backend(tmp_path):
    import dynamo_release_b200 as dvf
    from dynamo_release_b200 import patch

    ref_loader.install()
    VF = ref_loader.load("dynamo.vectorfield.VectorField")
    vc = ref_loader.load("dynamo.vectorfield.vector_calculus")
    X, V = o.hematopoiesis_like(1500, 30, seed=6)
    ad = make_pca_adata(X, V)
    patch.install()
    try:
        VF.VectorField(ad, basis="pca", M=80)
        vc.jacobian(ad, regulators=["g001", "g004"], effectors=["g002", "g003"])
        if ref_loader.available():
            try:
                cleanup = ref_loader.load("dynamo.data_io").cleanup
                ad.uns["pca_fit"], ad.uns["kinetics_heatmap"] = object(), {"x": 1}
                ad.uns["VecFld_pca"]["VecFld2D"] = object()          # what topography leaves behind (:316-319)
                cleanup(ad)
                assert ad.uns["pca_fit"] is None and "kinetics_heatmap" not in ad.uns
                assert "VecFld2D" not in ad.uns["VecFld_pca"]
            except ImportError:                                       # pragma: no cover
                pass
        vf_mem = ad.uns["VecFld_pca"]
        path = os.path.join(tmp_path, "adata_uns.npz")
        save_uns(path, ad.uns)
        ad2 = make_pca_adata(X, V)
        ad2.uns = load_uns(path)
        vf_disk = ad2.uns["VecFld_pca"]
        # the leaves that were None (grid, grid_V, dims) are gone, everything else is there with equal values
        assert set(vf_mem) - set(vf_disk) == {k for k, v in vf_mem.items() if v is None}
        for k in vf_disk:
            assert np.array_equal(np.asarray(vf_disk[k]), np.asarray(vf_mem[k])), k
        assert isinstance(vf_disk["method"], str) and isinstance(vf_disk["beta"], np.floating)
        # vecfld_from_adata / from_adata on the reloaded object: same field, bit for bit
        d, func = dvf.vecfld_from_adata(ad2, basis="pca")
        assert d is vf_disk
        xs = X[::7]
        assert np.array_equal(func(xs), dvf.vector_field_function(xs, vf_mem))
        svc = dvf.SvcVectorField()
        svc.from_adata(ad2, basis="pca")
        assert np.array_equal(svc.get_Jacobian()(xs), dvf.Jacobian_rkhs_gaussian(xs, vf_mem))
        assert np.array_equal(svc.compute_divergence(xs), dvf.compute_divergence(dvf.RkhsJacobian(vf_mem), xs))
        # the adata-level consumers run on the reloaded object as on the original one
        vc.divergence(ad2, basis="pca")
        vc.divergence(ad, basis="pca")
        assert np.array_equal(ad2.obs["divergence_pca"].values, ad.obs["divergence_pca"].values)
        vc.acceleration(ad2, basis="pca")
        vc.acceleration(ad, basis="pca")
        assert np.array_equal(ad2.obsm["acceleration_pca"], ad.obsm["acceleration_pca"])
        j_mem, j_disk = ad.uns["jacobian_pca"], ad2.uns["jacobian_pca"]
        assert np.array_equal(j_disk["jacobian"], j_mem["jacobian"]) and j_disk["jacobian_gene"].dtype == np.float32
        assert list(j_disk["regulators"]) == list(j_mem["regulators"]) and np.array_equal(j_disk["cell_idx"], j_mem["cell_idx"])
    finally:
        patch.uninstall()
