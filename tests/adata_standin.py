"""AnnData STAND-IN for the adata-level tests (`anndata` is not installed in this image).

Just enough of the AnnData surface for the UNMODIFIED reference wrappers -- dynamo/vectorfield/VectorField.py:21-276,
vector_calculus.py:199-339 (jacobian), :756-809 (acceleration), :812-868 (curvature), :952-1020 (divergence),
rank_vf.py:440-569 (rank_jacobian_genes), tools/utils.py:238-280 (create_layer) -- to run against either backend:
dict-backed obsm / uns / layers / varm, pandas obs / var, obs_names / var_names, n_obs / n_vars / shape, X, uns_keys(),
and `adata[:, mask]` giving a column subset with its own var / X / layers.  It is NOT a subclass of the stubbed
anndata.AnnData, so the reference's @monitor structure watcher (tools/_track.py:284-295) leaves it alone.
Real-anndata runs remain "unverified here".
"""
from __future__ import annotations

import copy

import numpy as np
import pandas as pd


class AnnDataStandIn:
    def __init__(self, X, obs=None, var=None, obsm=None, uns=None, layers=None, varm=None):
        self.X = np.asarray(X)
        n, g = self.X.shape
        self.obs = obs if obs is not None else pd.DataFrame(index=pd.Index(["cell_%d" % i for i in range(n)]))
        self.var = var if var is not None else pd.DataFrame(index=pd.Index(["gene_%d" % i for i in range(g)]))
        self.obsm = dict(obsm or {})
        self.uns = dict(uns or {})
        self.layers = dict(layers or {})
        self.varm = dict(varm or {})

    # ---- names / shapes ----
    @property
    def obs_names(self):
        return self.obs.index

    @property
    def var_names(self):
        return self.var.index

    @property
    def n_obs(self):
        return self.X.shape[0]

    @property
    def n_vars(self):
        return self.X.shape[1]

    @property
    def shape(self):
        return self.X.shape

    def uns_keys(self):
        return list(self.uns.keys())

    def obsm_keys(self):
        return list(self.obsm.keys())

    def copy(self):
        return copy.deepcopy(self)

    # ---- adata[:, genes] / adata[cells, :] (views are materialised: the wrappers only read from them) ----
    def __getitem__(self, key):
        if not isinstance(key, tuple):
            key = (key, slice(None))
        rows, cols = key
        ri = self._indexer(rows, self.obs.index)
        ci = self._indexer(cols, self.var.index)
        sub = AnnDataStandIn(
            self.X[ri][:, ci],
            obs=self.obs.iloc[ri].copy(),
            var=self.var.iloc[ci].copy(),
            obsm={k: np.asarray(v)[ri] for k, v in self.obsm.items()},
            uns=self.uns,
            layers={k: v[ri][:, ci] for k, v in self.layers.items()},
            varm={k: np.asarray(v)[ci] for k, v in self.varm.items()},
        )
        return sub

    @staticmethod
    def _indexer(sel, index):
        n = len(index)
        if isinstance(sel, slice):
            return np.arange(n)[sel]
        if isinstance(sel, pd.Series):
            sel = sel.values
        arr = np.asarray(sel)
        if arr.dtype == bool:
            return np.nonzero(arr)[0]
        if arr.dtype.kind in "iu":
            return arr.reshape(-1)
        return index.get_indexer(list(np.atleast_1d(arr)))   # names


def snapshot(adata):
    """Every array / scalar the wrappers wrote, flattened to {path: value} for key-by-key comparison."""
    out = {}

    def walk(prefix, v):
        if isinstance(v, dict):
            out[prefix + "/<keys>"] = sorted(map(str, v.keys()))
            for k, x in v.items():
                walk("%s/%s" % (prefix, k), x)
        else:
            out[prefix] = v

    for slot in ("obsm", "uns", "layers", "varm"):
        walk(slot, getattr(adata, slot))
    out["obs/<columns>"] = list(adata.obs.columns)
    for c in adata.obs.columns:
        out["obs/%s" % c] = adata.obs[c].values
    return out


def make_pca_adata(X_pca, V_pca, n_genes=40, seed=5):
    """Cells with a PCA embedding + velocity and an orthonormal loading matrix in uns['PCs'] (genes x PCs), the fields
    VectorField(basis='pca') / jacobian / acceleration / curvature read."""
    n, d = X_pca.shape
    rng = np.random.default_rng(seed)
    g = max(n_genes, d)
    Q, _ = np.linalg.qr(rng.normal(size=(g, d)))
    var = pd.DataFrame(index=pd.Index(["g%03d" % i for i in range(g)]))
    var["use_for_dynamics"] = True
    var["use_for_pca"] = True
    var["use_for_transition"] = True
    ad = AnnDataStandIn(np.zeros((n, g)), var=var, obsm={"X_pca": X_pca.copy(), "velocity_pca": V_pca.copy()},
                        uns={"PCs": np.ascontiguousarray(Q)})
    return ad
