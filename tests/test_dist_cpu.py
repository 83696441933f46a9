"""CPU (gloo, world_size 2): host-side logic of the sharded fit -- shard bounds, ragged row gather, and the
algebra the multi-GPU EM relies on: summing per-shard Gram/RHS blocks and scalars reproduces the single-shard
M-step of the oracle (reference scVectorField.py:468-486)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from dynamo_release_b200.dist import gather_rows, shard_bounds
from oracle import vfc_oracle as o


def test_shard_bounds_cover_everything():
    for n in (0, 1, 7, 1000, 1_000_003):
        for w in (1, 2, 3, 8):
            b = [shard_bounds(n, w, r) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            assert max(hi - lo for lo, hi in b) - min(hi - lo for lo, hi in b) <= 1


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        X, V = o.hematopoiesis_like(3001, 12, seed=4)
        V *= 0.05
        lo, hi = shard_bounds(len(X), world, rank)
        Xl, Vl = torch.from_numpy(X[lo:hi]), torch.from_numpy(V[lo:hi])
        # ragged gather restores the global row order on rank 0
        Xg, counts = gather_rows(Xl)
        if rank == 0:
            assert counts == [shard_bounds(len(X), world, r)[1] - shard_bounds(len(X), world, r)[0] for r in range(world)]
            assert np.array_equal(Xg.numpy(), X)
        # one EM iteration: per-shard partial sums + allreduce == single-shard oracle quantities
        _, _, _, ctrl, _, beta = o.prelude(X, V, 64)
        Kc = o.gauss_kernel(ctrl, ctrl, beta)
        D = X.shape[1]
        sigma2 = float(np.sum(V**2) / V.size)
        # E-step needs the GLOBAL min non-zero t1 (scVectorField.py:223): allreduce(MIN)
        r2l = np.sum(V[lo:hi] ** 2, 1)
        t1 = np.exp(-r2l / (2 * sigma2))
        mn = torch.tensor([t1[t1 != 0].min()])
        dist.all_reduce(mn, op=dist.ReduceOp.MIN)
        t1[t1 == 0] = mn.item()
        t2 = (2 * np.pi * sigma2) ** (D / 2) * (1 - 0.9) / (0.9 * 5)
        P = np.maximum(t1 / (t1 + t2), 1e-5)
        U = o.gauss_kernel(X[lo:hi], ctrl, beta)
        blk = np.concatenate([(U.T * P) @ U, (U.T * P) @ V[lo:hi]], 1)
        tb = torch.from_numpy(np.concatenate([blk.ravel(), [P.sum()]]))
        dist.all_reduce(tb)
        if rank == 0:
            Pf, _ = o.posterior(V, np.zeros_like(V), sigma2, 0.9, 5)
            Pf = np.maximum(Pf, 1e-5)
            lhs, rhs = o.m_step_system(o.gauss_kernel(X, ctrl, beta), Pf, V, Kc, 3.0, sigma2)
            M = ctrl.shape[0]
            got = tb.numpy()[:-1].reshape(M, M + D)
            assert np.allclose(got[:, :M] + 3.0 * sigma2 * Kc, lhs, rtol=1e-12, atol=0)
            assert np.allclose(got[:, M:], rhs, rtol=1e-11, atol=1e-14)
            assert abs(tb.numpy()[-1] - Pf.sum()) <= 1e-12 * Pf.sum()
        ret[rank] = "ok"
    finally:
        dist.destroy_process_group()


def test_sharded_m_step_algebra_gloo_world2():
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), ret), nprocs=world, join=True)
    assert dict(ret) == {0: "ok", 1: "ok"}


def _prelude_case(case):
    X, V = o.hematopoiesis_like(2503, 9, seed=11 + case)
    V *= 0.05
    if case >= 1:
        # duplicate rows (first occurrence must win), rows that tie on the first coordinate only, across shards
        X[40] = X[2000]
        X[1500] = X[7]
        X[2400] = X[7]
        X[100, 0] = X[1900, 0]
        X[101, 0] = X[1900, 0]
        X[102, :2] = X[1900, :2]
    if case == 2:
        X[:, 0] = np.round(X[:, 0], 1)          # heavy ties on the first coordinate
    return X, V


def _prelude_worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from dynamo_release_b200 import prelude
        from dynamo_release_b200.dist import sharded_prelude
        for case in range(3):
            X, V = _prelude_case(case)
            for vbs in (True, False):
                lo, hi = shard_bounds(len(X), world, rank)
                ctrl, idx, beta, n_total = sharded_prelude(torch.from_numpy(X[lo:hi].copy()), torch.from_numpy(V[lo:hi].copy()),
                                                           60, seed=3, velocity_based_sampling=vbs)
                _, _, _, ctrl1, idx1, beta1 = prelude.run(torch.from_numpy(X.copy()), torch.from_numpy(V.copy()), 60, seed=3,
                                                          velocity_based_sampling=vbs)
                assert n_total == len(X)
                assert np.array_equal(idx, idx1), (case, vbs)
                assert np.array_equal(ctrl.numpy(), ctrl1), (case, vbs)
                assert beta == beta1
                # and both equal the oracle's NumPy prelude (np.unique / Generator.choice themselves)
                if vbs:
                    _, _, _, ctrl_o, idx_o, beta_o = o.prelude(X, V, 60, seed=3)
                    assert np.array_equal(idx, idx_o) and np.array_equal(ctrl.numpy(), ctrl_o)
                    assert abs(beta - beta_o) <= 1e-13 * beta_o
        ret[rank] = "ok"
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_prelude_equals_single_process(world):
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_prelude_worker, args=(world, _free_port(), ret), nprocs=world, join=True)
    assert dict(ret) == {r: "ok" for r in range(world)}


def _ivp_worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import ivp_util
        from dynamo_release_b200 import integrate as it
        from dynamo_release_b200.dist import integrate_sharded

        # the NumPy oracle stands in for the device solver: this checks the sharding / gathering logic only
        it.solve_ivp_batch = ivp_util.oracle_batch(it.DenseSolution, it.IvpResult)
        z = ivp_util.load()
        f = it.RkhsField(ivp_util.field(z, "sink"))
        init = z["sink/init"]
        t = np.linspace(0, 3.0, 7)
        tt, Y = integrate_sharded(init, t, "forward", f, interpolation_num=int(z["interp"]), sampling="logspace",
                                  disable=True)
        if rank == 0:
            key = "sink/T3/forward/logspace/avg1"
            assert len(Y) == len(init)
            for i in range(len(Y)):
                assert np.allclose(Y[i], z[f"{key}/Y{i}"], rtol=0, atol=1e-10 * np.abs(z[f"{key}/Y{i}"]).max())
                assert np.allclose(tt[i], z[f"{key}/t{i}"], rtol=1e-12)
        else:
            lo, hi = shard_bounds(len(init), world, rank)
            assert len(Y) == hi - lo
        with pytest.raises(NotImplementedError):
            integrate_sharded(init, t, "forward", f, sampling="uniform_indices")
        ret[rank] = True
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_trajectory_integration_gloo(world):
    ret = mp.Manager().dict()
    mp.spawn(_ivp_worker, args=(world, _free_port(), ret), nprocs=world, join=True)
    assert all(ret.get(r) for r in range(world))
