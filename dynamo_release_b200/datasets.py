"""Synthetic 'hematopoiesis-shaped' workloads: a branching low-rank manifold in PCA space
with decaying per-PC variance, velocities from a smooth field plus noise and a fraction of outliers.  Host-side
data generation for bench.py / tests only -- not part of the hot path."""
from __future__ import annotations

import numpy as np


def hematopoiesis_like(n_cells, n_pcs=30, seed=0, outlier_frac=0.05, velocity_scale=1.0, manifold_seed=None):
    """`manifold_seed`: draw the embedding matrix A from its own stream, so that different `seed`s give different
    cells of the SAME manifold (the shards of a multi-GPU run)."""
    rng = np.random.default_rng(seed)
    t = rng.uniform(0, 1, n_cells)
    b = rng.choice(3, size=n_cells, p=[0.5, 0.3, 0.2])
    z = np.stack([t, t**2 * (b == 1), t**2 * (b == 2), np.sin(3 * t)], 1)
    dz = np.stack([np.ones_like(t), 2 * t * (b == 1), 2 * t * (b == 2), 3 * np.cos(3 * t)], 1)
    A = rng.normal(size=(4, n_pcs))
    if manifold_seed is not None:
        A = np.random.default_rng(manifold_seed).normal(size=(4, n_pcs))
    s = 3 * 0.85 ** np.arange(n_pcs)
    X = z @ A + s[None, :] * rng.normal(scale=0.05, size=(n_cells, n_pcs))
    V = (dz @ A) * (1 + 0.5 * np.sin(2 * np.pi * t))[:, None] + 0.02 * rng.normal(size=(n_cells, n_pcs))
    n_out = int(outlier_frac * n_cells)
    if n_out:
        rows = rng.choice(n_cells, n_out, replace=False)
        d = rng.normal(size=(n_out, n_pcs))
        V[rows] = 3 * d / np.linalg.norm(d, axis=1, keepdims=True)
    if velocity_scale != 1.0:
        V = V * velocity_scale
    return np.ascontiguousarray(X), np.ascontiguousarray(V)
