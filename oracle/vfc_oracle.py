"""TEST INFRASTRUCTURE ONLY -- CPU (NumPy/SciPy, float64) restatement of dynamo's SparseVFC hot path.

This file is the *oracle* ("port") for the B200 kernels.  It is imported only by tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.  The product package
(dynamo_release_b200/) never imports it and has no CPU fallback.

Parity status: PINNED against the reference itself.  oracle/gen_golden.py executes the unmodified
reference source (oracle/ref_loader.py) on seeded inputs in the build container and commits the
outputs under tests/golden/; tests/test_oracle_golden.py checks this restatement against those
vectors (bit-exact or <=1e-13, see the test).  The reference's own test-suite holds no numeric
vectors for this path, so the executed reference is the only pin there is.

Each function cites the reference lines it restates (paths relative to /root/reference/dynamo).
The restatement keeps the reference's *cost structure* on purpose (materialised U, the M x N
weighted temporary, per-cell Python loops for the Jacobian) because it doubles as the timed CPU
baseline.
"""
from __future__ import annotations

import numpy as np
from scipy.linalg.blas import dgemm
from scipy.spatial.distance import cdist


# --------------------------------------------------------------------------------------------
# kernel matrix, field evaluation
# --------------------------------------------------------------------------------------------
def gauss_kernel(x, y, beta, with_diff=False):
    """k(x_i, y_j) = exp(-beta |x_i - y_j|^2).   vectorfield/utils.py:180-212 (con_K).

    values-only path: scipy cdist 'sqeuclidean' (utils.py:194-197, flattened for a single row);
    with_diff path: broadcast differences diff[n, d, m] = x[n, d] - y[m, d] (utils.py:199-205)."""
    if not with_diff:
        sq = cdist(x, y, "sqeuclidean")
        if len(sq) == 1:
            sq = sq.flatten()
        return np.exp(-beta * sq)
    diff = x[:, :, None] - y.T[None, :, :]
    sq = np.squeeze(np.sum(diff**2, 1))
    return np.exp(-beta * sq), diff


def field_eval(x, vf, dim=None, ctrl_subset=None):
    """f(x) = K(x, X_ctrl) C.   vectorfield/utils.py:73-142 (vector_field_function)."""
    x = np.asarray(x)
    if x.ndim == 1:
        x = x[None, :]
    K = gauss_kernel(x, vf["X_ctrl"], vf["beta"])
    C = vf["C"]
    if ctrl_subset is not None:  # utils.py:128-132
        Cm = np.zeros_like(C)
        Cm[ctrl_subset, :] = C[ctrl_subset, :]
        C = Cm
    out = K.dot(C)
    if dim is not None:  # utils.py:136-140
        out = out[:, :dim] if np.isscalar(dim) else out[:, dim]
    return out


# --------------------------------------------------------------------------------------------
# prelude: valid rows, unique rows, control-point sampling, bandwidth
# --------------------------------------------------------------------------------------------
def knn_bandwidth(ctrl):
    """h = sqrt(2) * mean(kNN distance) / 1.5 with k+1 = max(2, int(0.2 M)) neighbours.

    vectorfield/scVectorField.py:121-135 -> tools/connectivity.py:484-577 (sklearn exact tree:
    kd_tree for D<=10, ball_tree above; both exact, so the method only changes rounding of ties)."""
    from sklearn.neighbors import NearestNeighbors

    n, d = ctrl.shape
    k = max(2, int(0.2 * n)) - 1
    algo = "ball_tree" if d > 10 else "kd_tree"
    nn = NearestNeighbors(n_neighbors=k + 1, metric="euclidean", algorithm=algo, n_jobs=-1).fit(ctrl)
    dist, _ = nn.kneighbors(ctrl)
    return np.sqrt(2) * (np.mean(dist[:, 1:]) / 1.5)


def velocity_weighted_choice(Vu, M, seed):
    """tools/sampling.py:231-246 (sample_by_velocity): Generator.choice without replacement, p ~ |v|."""
    rng = np.random.default_rng(seed)
    speed = np.linalg.norm(Vu, axis=1)
    return rng.choice(np.arange(len(Vu)), size=M, p=speed / np.sum(speed), replace=False)


def prelude(X, Y, M, seed=0, beta=None, velocity_based_sampling=True):
    """vectorfield/scVectorField.py:396-415.  Returns valid_ind, Xv, Yv, ctrl_pts, idx, beta."""
    valid = np.where(np.isfinite(Y.sum(1)))[0]
    Xv, Yv = X[valid], Y[valid]
    uniq, first = np.unique(Xv, axis=0, return_index=True)
    M = min(M, uniq.shape[0])
    if velocity_based_sampling:
        idx = velocity_weighted_choice(Yv[first], M, seed)
    else:
        idx = np.random.RandomState(seed=seed).permutation(uniq.shape[0])[:M]
    ctrl = uniq[idx, :]
    if beta is None:
        h = knn_bandwidth(ctrl)
        beta = 1 / h**2
    return valid, Xv, Yv, ctrl, idx, beta


# --------------------------------------------------------------------------------------------
# EM pieces
# --------------------------------------------------------------------------------------------
def posterior(Y, V, sigma2, gamma, a):
    """E-step.  vectorfield/scVectorField.py:197-227 (get_P), Gaussian-kernel branch."""
    D = Y.shape[1]
    r2 = np.sum((Y - V) ** 2, 1)
    t1 = np.exp(-r2 / (2 * sigma2))
    t2 = (2 * np.pi * sigma2) ** (D / 2) * (1 - gamma) / (gamma * a)
    t1[t1 == 0] = np.min(t1[t1 != 0])  # :223 -- global min-nonzero replaces exact zeros
    P = t1 / (t1 + t2)
    E = P.T.dot(np.sum((Y - V) ** 2, 1)) / (2 * sigma2) + np.sum(P) * np.log(sigma2) * D / 2
    return P[:, None], E


def normal_eq_solve(A, B):
    """tools/utils.py:2948-2981 (linear_least_squares, 'drouin'): solve(A^T A, A^T B) via dgemm + gesv."""
    A = np.asarray(A, order="c")
    AtA = dgemm(alpha=1.0, a=A.T, b=A.T, trans_b=True)
    return np.linalg.solve(AtA, dgemm(alpha=1.0, a=A.T, b=B))


def m_step_system(U, P, Y, Kc, lam, sigma2):
    """lhs = U^T diag(P) U + lam sigma2 K ; rhs = U^T diag(P) Y.   scVectorField.py:474-476.
    The M x N weighted temporary is formed exactly as the reference does."""
    M = U.shape[1]
    UP = U.T * np.tile(P.T, (M, 1))
    return UP.dot(U) + lam * sigma2 * Kc, UP.dot(Y)


def sparse_vfc(X, Y, M=100, a=5, beta=None, ecr=1e-5, gamma=0.9, lambda_=3, minP=1e-5, MaxIter=500,
               theta=0.75, velocity_based_sampling=True, seed=0, trace=None):
    """SparseVFC EM fit, Gaussian kernel.   vectorfield/scVectorField.py:313-542.

    `trace`, if a list, receives one dict per iteration with the quantities a step-level
    ("teacher-forced") parity test needs: inputs of the iteration and everything it produced."""
    X_in, Y_in = X.copy(), Y.copy()
    valid, X, Y, ctrl, idx, beta = prelude(X, Y, M, seed, beta, velocity_based_sampling)
    N, D = Y.shape
    Kc = gauss_kernel(ctrl, ctrl, beta)
    U = gauss_kernel(X, ctrl, beta)
    M = ctrl.shape[0]

    V = np.zeros((N, D))
    C = np.zeros((M, D))
    i, tecr, E = 0, 1, 1
    sigma2 = sum(sum((Y - V) ** 2)) / (N * D)  # builtin sum on purpose (:444) -- sequential rounding
    sigma2 = 1e-7 if sigma2 < 1e-8 else sigma2
    tecr_vec = np.full(MaxIter, np.nan)
    E_vec = np.full(MaxIter, np.nan)
    P = None
    while i < MaxIter and tecr > ecr and sigma2 > 1e-8:
        E_old = E
        step = {"sigma2_in": sigma2, "gamma_in": gamma, "C_in": C, "V_in": V} if trace is not None else None
        P, E = posterior(Y, V, sigma2, gamma, a)
        E = E + lambda_ / 2 * np.trace(C.T.dot(Kc).dot(C))
        E_vec[i] = E
        tecr = abs((E - E_old) / E)
        tecr_vec[i] = tecr

        P = np.maximum(P, minP)
        lhs, rhs = m_step_system(U, P, Y, Kc, lambda_, sigma2)
        C = normal_eq_solve(lhs, rhs)

        V = U.dot(C)
        Sp = sum(P)
        sigma2 = (sum(P.T.dot(np.sum((Y - V) ** 2, 1))) / np.dot(Sp, D))[0]
        numcorr = len(np.where(P > theta)[0])
        gamma = min(max(numcorr / X.shape[0], 0.05), 0.95)
        if step is not None:
            step.update(P=P, E=E, tecr=tecr, lhs=lhs, rhs=rhs, C=C, V=V, sigma2=sigma2, gamma=gamma)
            trace.append(step)
        i += 1
    if i == 0:
        raise Exception(
            "please check your input parameters, "
            f"tecr: {tecr}, ecr {ecr} and sigma2 {sigma2},"
            f"tecr must larger than ecr and sigma2 must larger than 1e-8"
        )
    return {
        "X": X_in, "valid_ind": valid, "X_ctrl": ctrl, "ctrl_idx": idx, "Y": Y_in, "beta": beta,
        "V": V, "C": C, "P": P, "VFCIndex": np.where(P > theta)[0], "sigma2": sigma2,
        "grid": None, "grid_V": None, "iteration": i - 1, "tecr_traj": tecr_vec[:i], "E_traj": E_vec[:i],
    }


def train(X, V, **params):
    """SvcVectorField(X, V, None, normalize=False).train().   scVectorField.py:997-1046:
    fit, then re-evaluate V = f(X) on ALL rows (:1043)."""
    vf = sparse_vfc(X, V, **params)
    vf["V"] = field_eval(X, vf)
    vf["normalize"] = False
    return vf


# --------------------------------------------------------------------------------------------
# differential geometry
# --------------------------------------------------------------------------------------------
def jacobian(x, vf):
    """J[:, :, n] = -2 beta (C^T * k_n) (x_n - Xc)^T.   vectorfield/utils.py:351-379, default
    (per-cell loop) branch; output (D, D, n), (D, D) for a single point."""
    Xc, beta, C = vf["X_ctrl"], vf["beta"], vf["C"]
    if x.ndim == 1:
        k, diff = gauss_kernel(x[None, :], Xc, beta, with_diff=True)
        return -2 * beta * ((C.T * k) @ diff[0].T)
    n, d = x.shape
    J = np.zeros((d, d, n))
    for i, xi in enumerate(x):
        k, diff = gauss_kernel(xi[None, :], Xc, beta, with_diff=True)
        J[:, :, i] = (C.T * k) @ diff[0].T
    return -2 * beta * J


def divergence(x, vf, batch=1000):
    """tr J per cell in batches.   vectorfield/utils.py:699-725."""
    n = len(x)
    out = np.zeros(n)
    for s in range(0, n, batch):
        out[s:s + batch] = np.trace(jacobian(x[s:s + batch], vf))
    return out


def acceleration(x, vf, Js=None):
    """a_n = J_n v_n, |a_n|.   vectorfield/utils.py:770-793.  Returns v, J, |a|, a."""
    n = len(x)
    v = field_eval(x, vf)
    J = jacobian(x, vf) if Js is None else Js
    a = np.zeros((n, x.shape[1]))
    an = np.zeros(n)
    for i in range(n):
        a[i] = J[:, :, i].dot(v[i][:, None]).flatten()
        an[i] = np.linalg.norm(a[i])
    return v, J, an, a


def curvature(x, vf, formula=2, Js=None):
    """vectorfield/utils.py:797-821 with curvature_method2 (:752-757) / method1 (:743-749)."""
    n = len(x)
    v, _, _, a = acceleration(x, vf, Js=Js)
    kn = np.zeros(n)
    kv = np.zeros((n, x.shape[1])) if formula == 2 else None
    for i in range(n):
        if formula == 1:
            kn[i] = np.linalg.norm(np.outer(v[i][:, None], a[i])) / np.linalg.norm(v[i][:, None]) ** 3
        else:
            kv[i] = (np.multiply(a[i], np.dot(v[i], v[i])) - np.multiply(v[i], np.dot(v[i], a[i]))) \
                / np.linalg.norm(v[i]) ** 4
            kn[i] = np.linalg.norm(kv[i])
    return kn, kv


def gene_jacobian(Js, Qe, Qr):
    """Jhat[:, :, n] = Qe J_n Qr^T stored as float32.   vectorfield/utils.py:544-554."""
    Js = np.atleast_3d(Js)
    Qe, Qr = np.atleast_2d(Qe), np.atleast_2d(Qr)
    out = np.zeros((Qe.shape[0], Qr.shape[0], Js.shape[2]), dtype=np.float32)
    for i in range(Js.shape[2]):
        out[:, :, i] = Qe @ Js[:, :, i] @ Qr.T
    return out


def gene_jacobian_element(Js, qe, qr):
    """q_e J_n q_r per cell (float64).   vectorfield/utils.py:430-453."""
    Js = np.atleast_3d(Js)
    out = np.zeros(Js.shape[2])
    for i in range(Js.shape[2]):
        out[i] = qe @ Js[:, :, i] @ qr
    return out


def to_gene_space(V, Q):
    """vectorfield/utils.py:311-326."""
    return V @ Q.T


# --------------------------------------------------------------------------------------------
# per-cell post-processing
# --------------------------------------------------------------------------------------------
def jacobian_det(Js):
    """vectorfield/vector_calculus.py:329: np.linalg.det of every cell's Jacobian."""
    Js = np.atleast_3d(Js)
    return np.array([np.linalg.det(Js[:, :, i]) for i in range(Js.shape[2])])


def vector_angle(v1, v2):
    """vectorfield/utils.py:1148-1173 (angle / unit_vector)."""
    n1, n2 = np.linalg.norm(v1), np.linalg.norm(v2)
    if n1 == 0 or n2 == 0:
        return np.nan
    u1, u2 = v1 / n1, v2 / n2
    minor = np.linalg.det(np.stack((u1[-2:], u2[-2:])))
    sign = 1 if minor == 0 else -np.sign(minor)
    return sign * np.arccos(min(max(np.dot(u1, u2), -1.0), 1.0))


def cell_angles(U, V):
    """vectorfield/VectorField.py:255-259: angle between observed and fitted velocity, per cell."""
    return np.array([vector_angle(u.astype("float64"), v.astype("float64")) for u, v in zip(U, V)])


def speed(V):
    """vectorfield/vector_calculus.py:192."""
    return np.array([np.linalg.norm(i) for i in V])


# --------------------------------------------------------------------------------------------
# second derivatives and low-dimensional calculus helpers
# --------------------------------------------------------------------------------------------
def hessian(x, vf):
    """vectorfield/utils.py:589-614, one point x (1-D): H[i, l, m]."""
    x = np.atleast_2d(x)
    K, Dm = gauss_kernel(x, vf["X_ctrl"], vf["beta"], with_diff=True)     # K (1, M), Dm (1, D, M)
    K = np.atleast_2d(K) * vf["C"].T                                      # (D, M): k_j C_ji
    Dm = Dm.T                                                             # (M, D, 1)
    Dm = np.eye(x.shape[1]) - 2 * vf["beta"] * Dm @ np.transpose(Dm, axes=(0, 2, 1))
    return -2 * vf["beta"] * np.einsum("ij, jlm -> ilm", K, Dm)


def hessian_cells(X, vf):
    """vectorfield/vector_calculus.py:413-417: Hs[:, :, :, c] = Hessian(X[c])."""
    Hs = np.zeros([X.shape[1], X.shape[1], X.shape[1], X.shape[0]])
    for c in range(X.shape[0]):
        Hs[:, :, :, c] = hessian(X[c], vf)
    return Hs


def laplacian_of(H):
    """vectorfield/utils.py:665-687."""
    if H.ndim == 4:
        L = np.zeros([H.shape[2], H.shape[3]])
        for c in range(H.shape[3]):
            for k in range(L.shape[0]):
                L[k, c] = np.diag(H[:, :, k, c]).sum()
    else:
        L = np.zeros([H.shape[2], 1])
        for k in range(L.shape[0]):
            L[k, 0] = np.diag(H[:, :, k]).sum()
    return L


def curl_cells(X, vf):
    """vectorfield/utils.py:873-905 with the analytical Jacobian."""
    n = len(X)
    out = np.zeros(n) if X.shape[1] == 2 else np.zeros((n, 3, 3))
    for i in range(n):
        J = jacobian(X[i], vf)
        out[i] = (J[1, 0] - J[0, 1]) if X.shape[1] == 2 else np.array([J[2, 1] - J[1, 2], J[0, 2] - J[2, 0], J[1, 0] - J[0, 1]])
    return out


def torsion_cells(X, vf):
    """vectorfield/utils.py:760-766, 823-841."""
    n = len(X)
    tor = np.zeros((n, 3, 3))
    v = field_eval(X, vf)
    for i in range(n):
        J = jacobian(X[i], vf)
        a = J.dot(v[i][:, None]).flatten()
        va = np.outer(v[i][:, None], a)
        tor[i] = va.dot(J.dot(a)) / np.linalg.norm(va) ** 2
    return tor


def sensitivity_cells(X, vf):
    """vectorfield/utils.py:844-870."""
    J = jacobian(X, vf)
    S = np.zeros_like(J)
    eye = np.eye(J.shape[0])
    for i in range(J.shape[2]):
        s = np.linalg.inv(eye - J[:, :, i])
        S[:, :, i] = s.dot(np.diag(1 / np.diag(s)))
    return S


# --------------------------------------------------------------------------------------------
# normalisation (scVectorField.py:64-96, 138-182)
# --------------------------------------------------------------------------------------------
def normalise(X, V, fix_velocity=True):
    Y = X + V
    xm, ym = np.mean(X, axis=0), np.mean(Y, axis=0)
    x, y = X - xm[None, :], Y - ym[None, :]
    xs, ys = np.sqrt(np.mean(x**2, axis=0))[None, :], np.sqrt(np.mean(y**2, axis=0))[None, :]
    Xn, Yn = x / xs, y / ys
    return Xn, (V if fix_velocity else Yn - Xn), {"xm": xm, "ym": ym, "xscale": xs, "yscale": ys,
                                                  "fix_velocity": fix_velocity}


# --------------------------------------------------------------------------------------------
# synthetic workloads -- shared by tests, smoke and bench
# --------------------------------------------------------------------------------------------
def hematopoiesis_like(n_cells, n_pcs=30, seed=0, outlier_frac=0.05):
    """Branching low-rank manifold in PCA space with decaying per-PC variance; velocities from a
    smooth field plus noise; a fraction of rows replaced by random-direction outliers."""
    rng = np.random.default_rng(seed)
    t = rng.uniform(0, 1, n_cells)
    b = rng.choice(3, size=n_cells, p=[0.5, 0.3, 0.2])
    z = np.stack([t, t**2 * (b == 1), t**2 * (b == 2), np.sin(3 * t)], 1)
    dz = np.stack([np.ones_like(t), 2 * t * (b == 1), 2 * t * (b == 2), 3 * np.cos(3 * t)], 1)
    A = rng.normal(size=(4, n_pcs))
    s = 3 * 0.85 ** np.arange(n_pcs)
    X = z @ A + s[None, :] * rng.normal(scale=0.05, size=(n_cells, n_pcs))
    V = (dz @ A) * (1 + 0.5 * np.sin(2 * np.pi * t))[:, None] + 0.02 * rng.normal(size=(n_cells, n_pcs))
    n_out = int(outlier_frac * n_cells)
    if n_out:
        rows = rng.choice(n_cells, n_out, replace=False)
        d = rng.normal(size=(n_out, n_pcs))
        V[rows] = 3 * d / np.linalg.norm(d, axis=1, keepdims=True)
    return np.ascontiguousarray(X), np.ascontiguousarray(V)
