"""Ranking of regulators / effectors / interactions from the gene-space Jacobian: the consumer of
`transform_jacobian`, reference vectorfield/rank_vf.py:440-569 (`rank_jacobian_genes`) with the table helpers of
tools/utils.py:878-982.

`rank_jacobian_genes(adata, ...)` is the drop-in (same signature; works on the stored float32 `jacobian_gene`, means in
NumPy's own float32 summation order, rankings exact).  `rank_jacobian_genes_fused` is the scalable extension:
the reference first materialises `jacobian_gene` (Ge, Gr, n) float32 and then reduces it over cells; at 2000 x 2000
genes that array is 16 MB per cell.  There the reduction is fused behind the transform (kernels.gene_transform_reduce:
per-cell float32 values exactly as the reference stores them, summed in float64 in a fixed order), the sorts run on
the device, and only the rank tables come back.  Inputs are the PCA-space Jacobian `Js` (D, D, n) and the loading rows
of the effector / regulator genes -- what `dyn.vf.jacobian` holds before it calls the transform
(vector_calculus.py:300-313).
"""
from __future__ import annotations

import numpy as np
import pandas as pd
import torch

from . import kernels
from .device import dev, host


def jacobian_gene_means(Js, Qe, Qr, groups=None, abs=False, normalize=False, switch=False):
    """Per-group means over cells of the (transformed) gene-space Jacobian: rank_vf.py:483-508.
    Returns (list of group keys in the reference's dict order, float64 CUDA tensor (n_groups, Ge, Gr))."""
    Jd = dev(Js)
    if Jd.dim() == 2:
        Jd = Jd[:, :, None]
    n = Jd.shape[2]
    Qe, Qr = np.atleast_2d(np.asarray(Qe, dtype=np.float64)), np.atleast_2d(np.asarray(Qr, dtype=np.float64))
    if groups is None:
        keys, lab = ["all"], np.zeros(n, dtype=np.int32)
    else:
        grps = np.asarray(groups)
        if grps.shape[0] != n:
            raise Exception("The group information you provided does not match the number of cells.")
        # average_jacobian_by_group (utils.py:557-582) fills its dict in order of first appearance
        uniq, first, inv = np.unique(grps, return_index=True, return_inverse=True)
        order = np.argsort(first, kind="stable")
        rank_of = np.empty(len(uniq), dtype=np.int32)
        rank_of[order] = np.arange(len(uniq), dtype=np.int32)
        keys, lab = [uniq[i] for i in order], rank_of[inv.reshape(-1)]
    cnt = np.bincount(lab, minlength=len(keys)).astype(np.float64)
    jmax = None
    if normalize:
        jmax = kernels.gene_transform_reduce(Jd, Qe, Qr, op="max", take_abs=abs).to(torch.float32)
    sums = kernels.gene_transform_reduce(Jd, Qe, Qr, op="switch" if switch else "sum", take_abs=abs, jmax=jmax,
                                         group=lab, n_groups=len(keys))
    return keys, sums / torch.as_tensor(np.maximum(cnt, 1.0), device=sums.device)[:, None, None]


def _sorted_rows(mat: torch.Tensor, order: int):
    """np.argsort(row)[::order] for every row, on the device; returns (index, sorted values) as numpy."""
    idx = torch.argsort(mat, dim=-1)
    if order == -1:
        idx = idx.flip(-1)
    return host(idx.to(torch.int32)), host(torch.gather(mat, -1, idx))


def pairwise_leaves(n: int) -> np.ndarray:
    """NumPy's pairwise-summation tree for n contiguous elements (numpy/_core/src/umath/loops_utils.h) as a program for
    the device: rows (first element, length, additions that close right after this leaf), leaves left to right."""
    out = []

    def rec(lo, m):
        if m <= 128:
            out.append([lo, m, 0])
            return
        m2 = m // 2
        m2 -= m2 % 8
        rec(lo, m2)
        rec(lo + m2, m - m2)
        out[-1][2] += 1

    rec(0, int(n))
    return np.asarray(out, dtype=np.int32)


def _cell_means_numpy_order(J: torch.Tensor, labels=None):
    """Means over the cell axis of the materialised gene-space Jacobian J (Ge, Gr, n) float32 on the device, in the
    float32 summation order NumPy uses for the same statement of the reference:
      labels None -> {"all": np.mean(J, axis=2)}                         (rank_vf.py:503-504; pairwise float32 sum)
      labels      -> average_jacobian_by_group(J, labels)                 (utils.py:557-582; running float32 sums)
    Returns (keys in the reference's dict order, float32 CUDA tensor (n_groups, Ge, Gr))."""
    from . import _lib
    from .device import lib, stream_ptr

    Ge, Gr, n = J.shape
    J = J.contiguous()
    R = Ge * Gr
    if labels is None:
        leaves = torch.as_tensor(pairwise_leaves(n), device=J.device).contiguous()
        out = torch.empty((1, Ge, Gr), dtype=torch.float32, device=J.device)
        _lib.check(lib().dvf_f32_row_means_pairwise(J.data_ptr(), R, n, leaves.data_ptr(), leaves.shape[0],
                                                    out.data_ptr(), stream_ptr()))
        return ["all"], out
    grps = np.asarray(labels)
    uniq, first, inv = np.unique(grps, return_index=True, return_inverse=True)
    order = np.argsort(first, kind="stable")                    # dict insertion order = first appearance
    rank_of = np.empty(len(uniq), dtype=np.int32)
    rank_of[order] = np.arange(len(uniq), dtype=np.int32)
    lab = torch.as_tensor(rank_of[inv.reshape(-1)].astype(np.int32), device=J.device)
    cnt = torch.as_tensor(np.bincount(rank_of[inv.reshape(-1)], minlength=len(uniq)).astype(np.float32), device=J.device)
    out = torch.empty((len(uniq), Ge, Gr), dtype=torch.float32, device=J.device)
    _lib.check(lib().dvf_f32_group_means_seq(J.data_ptr(), R, n, lab.data_ptr(), len(uniq), cnt.data_ptr(), out.data_ptr(),
                                             stream_ptr()))
    return [uniq[i] for i in order], out


def rank_jacobian_genes(adata, groups=None, jkey="jacobian_pca", abs=False, mode="full reg", exclude_diagonal=False,
                        normalize=False, return_df=False, **kwargs):
    """Drop-in for dynamo.vectorfield.rank_vf.rank_jacobian_genes (reference rank_vf.py:440-569): same signature, same
    tables.  The O(Ge Gr n) part -- |.|, normalisation, the `switch` product and the means over cells of the stored
    float32 `jacobian_gene` -- runs on the device in the arithmetic (float32) and the summation order NumPy uses for the
    reference's statements, so the means are bit-identical and the rankings equal exactly; the O(Ge Gr log) tail (nanmean
    over genes, argsort, table assembly: tools/utils.py:878-982) is the reference's own few NumPy statements on the host."""
    J_dict = adata.uns[jkey]
    Jh = np.asarray(J_dict["jacobian_gene"])
    if Jh.dtype != np.float32:
        raise NotImplementedError("rank_jacobian_genes on the B200 backend serves the float32 `jacobian_gene` that "
                                  "dyn.vf.jacobian(basis='pca') stores; got %s" % Jh.dtype)
    if Jh.ndim != 3:
        raise ValueError("jacobian_gene must be (n_effectors, n_regulators, n_cells)")
    J = dev(Jh, torch.float32)
    if abs:
        J = torch.abs(J)
    if normalize:
        Jmax = torch.amax(torch.abs(J), dim=2)                    # exact whatever the order
        J = J / Jmax[:, :, None]                                 # J[:, :, i] /= Jmax for every cell (float32 division)
    if mode == "switch":
        Jt = J.transpose(0, 1)
        J = (J * Jt) * (torch.sign(J) == -1) * (torch.sign(Jt) == -1)
    labels = None
    if groups is not None:
        if type(groups) is str and groups in adata.obs.keys():
            grps = np.array(adata.obs[groups])
        elif hasattr(groups, "__len__") and not isinstance(groups, str):
            grps = np.array(groups)
        else:
            raise Exception(f"The group information {groups} you provided is not in your adata object.")
        labels = grps[J_dict["cell_idx"]]
    keys, means = _cell_means_numpy_order(J, labels)
    J_mean = {k: m for k, m in zip(keys, host(means))}           # {group: (Ge, Gr) float32}, the reference's dict

    # ---- from here on the reference's own statements (rank_vf.py:516-566) ----
    eff = np.array([x for x in J_dict["effectors"]])
    reg = np.array([x for x in J_dict["regulators"]])
    rank_dict = {}
    ov = kwargs.pop("output_values", True)
    if mode in ["full reg", "full_reg"]:
        for k, Jk in J_mean.items():
            rank_dict[k] = _table_top_genes(Jk, eff, reg, n_top_genes=None, output_values=ov, **kwargs)
    elif mode in ["full eff", "full_eff"]:
        for k, Jk in J_mean.items():
            rank_dict[k] = _table_top_genes(Jk.T, reg, eff, n_top_genes=None, output_values=ov, **kwargs)
    elif mode == "reg":
        for k, Jk in J_mean.items():
            if exclude_diagonal:
                for i, ef in enumerate(eff):
                    ii = np.where(reg == ef)[0]
                    if len(ii) > 0:
                        Jk[i, ii] = np.nan
            j = np.nanmean(Jk, axis=0)
            if ov:
                rank_dict[k], rank_dict[k + "_values"] = _list_top_genes(j, reg, None, return_sorted_array=True, **kwargs)
            else:
                rank_dict[k] = _list_top_genes(j, reg, None, **kwargs)
        rank_dict = pd.DataFrame(data=rank_dict)
    elif mode == "eff":
        for k, Jk in J_mean.items():
            if exclude_diagonal:
                for i, re in enumerate(reg):
                    ii = np.where(eff == re)[0]
                    if len(ii) > 0:
                        Jk[ii, i] = np.nan
            j = np.nanmean(Jk, axis=1)
            if ov:
                rank_dict[k], rank_dict[k + "_values"] = _list_top_genes(j, eff, None, return_sorted_array=True, **kwargs)
            else:
                rank_dict[k] = _list_top_genes(j, eff, None, **kwargs)
        rank_dict = pd.DataFrame(data=rank_dict)
    elif mode in ["int", "switch"]:
        for k, Jk in J_mean.items():
            ints, vals = _list_top_interactions(Jk, eff, reg, **kwargs)
            rank_dict[k] = []
            if ov:
                rank_dict[k + "_values"] = []
            for ind, int_val in enumerate(ints):
                if not (exclude_diagonal and int_val[0] == int_val[1]):
                    rank_dict[k].append(int_val[0] + " - " + int_val[1])
                    if ov:
                        rank_dict[k + "_values"].append(vals[ind])
        rank_dict = pd.DataFrame(data=rank_dict)
    else:
        raise ValueError(f"No such mode as {mode}.")
    if return_df:
        return rank_dict
    adata.uns[mode] = rank_dict


def _list_top_genes(arr, gene_names, n_top_genes=30, order=-1, return_sorted_array=False):
    """tools/utils.py:896-923."""
    imax = np.argsort(arr)[::order]
    if return_sorted_array:
        return gene_names[imax][:n_top_genes], arr[imax][:n_top_genes]
    return gene_names[imax][:n_top_genes]


def _list_top_interactions(mat, row_names, column_names, order=-1):
    """tools/utils.py:878-893, 926-948."""
    isort = np.argsort(mat, axis=None)[::order]
    rows, cols = isort // mat.shape[1], isort % mat.shape[1]
    ints = [[row_names[r], column_names[c]] for r, c in zip(rows, cols)]
    return ints, np.array([mat[r, c] for r, c in zip(rows, cols)])


def _table_top_genes(arrs, item_names, gene_names, return_df=True, output_values=False, **kwargs):
    """tools/utils.py:951-982."""
    table = {}
    for i, item in enumerate(item_names):
        if output_values:
            table[item], table[item + "_values"] = _list_top_genes(arrs[i], gene_names, return_sorted_array=True, **kwargs)
        else:
            table[item] = _list_top_genes(arrs[i], gene_names, **kwargs)
    return pd.DataFrame(data=table) if return_df else table


def rank_jacobian_genes_fused(Js, Qe, Qr, effectors, regulators, groups=None, abs=False, mode="full reg",
                              exclude_diagonal=False, normalize=False, output_values=True, n_top_genes=None, order=-1,
                              n_top_interactions=None):
    """EXTENSION (not a reference name): the tables of rank_jacobian_genes(..., return_df=True) from the PCA-space
    Jacobian and the loadings WITHOUT a stored `jacobian_gene` -- the (Ge, Gr, n) array (3.2 TB at config 4) is never
    materialised; per-cell float32 values as the reference stores them, reduced in float64 in a fixed order.  Because
    the reference accumulates those float32 values in float32, entries whose means tie to ~1e-7 relative can swap places
    against it; `rank_jacobian_genes` above is the exact drop-in.

    mode 'full reg' / 'full eff': dict {group: DataFrame}, one column per effector (regulator) listing ALL regulators
    (effectors) in rank order plus `<name>_values` columns; 'reg' / 'eff': DataFrame with one column of genes (and
    `_values`) per group; 'int' / 'switch': DataFrame of "effector - regulator" strings (and `_values`) per group.
    Like the reference (which passes n_top_genes=None to its table helpers) every list is complete unless
    `n_top_genes` / `n_top_interactions` (extensions) cut it."""
    eff = np.array([x for x in effectors])
    reg = np.array([x for x in regulators])
    if mode not in ("full reg", "full_reg", "full eff", "full_eff", "reg", "eff", "int", "switch"):
        raise ValueError(f"No such mode as {mode}.")
    keys, Jm = jacobian_gene_means(Js, Qe, Qr, groups=groups, abs=abs, normalize=normalize, switch=(mode == "switch"))
    ov = output_values
    rank_dict = {}
    if mode in ("full reg", "full_reg", "full eff", "full_eff"):
        by_reg = mode in ("full reg", "full_reg")
        items, genes = (eff, reg) if by_reg else (reg, eff)
        idx, val = _sorted_rows(Jm if by_reg else Jm.transpose(1, 2).contiguous(), order)
        for gi, k in enumerate(keys):
            table = {}
            for i, item in enumerate(items):
                table[item] = genes[idx[gi, i]]
                if ov:
                    table[item + "_values"] = val[gi, i]
            rank_dict[k] = pd.DataFrame(data=table)
        return rank_dict
    if mode in ("reg", "eff"):
        J = Jm.clone()
        if exclude_diagonal:
            same = torch.as_tensor(eff[:, None] == reg[None, :], device=J.device)
            J[:, same] = float("nan")
        j = torch.nanmean(J, dim=1 if mode == "reg" else 2)
        names = reg if mode == "reg" else eff
        idx, val = _sorted_rows(j, order)
        for gi, k in enumerate(keys):
            rank_dict[k] = names[idx[gi]][:n_top_genes]
            if ov:
                rank_dict[str(k) + "_values"] = val[gi][:n_top_genes]
        return pd.DataFrame(data=rank_dict)
    # 'int' / 'switch': every (effector, regulator) pair, sorted by the mean
    G, Ge, Gr = Jm.shape
    idx, val = _sorted_rows(Jm.reshape(G, Ge * Gr), order)
    top = Ge * Gr if n_top_interactions is None else min(int(n_top_interactions), Ge * Gr)
    for gi, k in enumerate(keys):
        e_i, r_i = idx[gi, :top] // Gr, idx[gi, :top] % Gr
        keep = np.ones(top, dtype=bool)
        if exclude_diagonal:
            keep = eff[e_i] != reg[r_i]
        rank_dict[k] = [a + " - " + b for a, b in zip(eff[e_i][keep], reg[r_i][keep])]
        if ov:
            rank_dict[str(k) + "_values"] = list(val[gi, :top][keep])
    return pd.DataFrame(data=rank_dict)
