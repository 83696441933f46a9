"""Ranking of regulators / effectors / interactions from the gene-space Jacobian: the consumer of
`transform_jacobian`, reference vectorfield/rank_vf.py:440-569 (`rank_jacobian_genes`) with the table helpers of
tools/utils.py:878-982.

The reference first materialises `jacobian_gene` (Ge, Gr, n) float32 and then reduces it over cells; at 2000 x 2000
genes that array is 16 MB per cell.  Here the reduction is fused behind the transform (kernels.gene_transform_reduce:
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


def rank_jacobian_genes(Js, Qe, Qr, effectors, regulators, groups=None, abs=False, mode="full reg",
                        exclude_diagonal=False, normalize=False, output_values=True, n_top_genes=None, order=-1,
                        n_top_interactions=None):
    """The tables of reference rank_jacobian_genes(..., return_df=True) (rank_vf.py:440-569), from the PCA-space
    Jacobian and the loadings instead of a stored `jacobian_gene`.

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
