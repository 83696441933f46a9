"""Multi-GPU SparseVFC: cells are sharded over ranks (one process per GPU); control points, C and the scalar EM
state are replicated; per EM iteration the only traffic is one allreduce of the [Gram | RHS | scalars] block
plus two scalar allreduces (DESIGN.md "multi-GPU").  torch.distributed is used for rendezvous and for moving
the prelude's inputs; the EM loop's collectives are NCCL calls issued by libdvf_b200 on the compute stream."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch
import torch.distributed as dist

from . import _lib, prelude
from .device import dev, lib
from .sparsevfc import fit_em


def shard_bounds(n_total: int, world: int, rank: int):
    """Contiguous cell slice [lo, hi) of `rank`."""
    base, rem = divmod(int(n_total), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


class Comm:
    """NCCL communicator owned by libdvf_b200, bootstrapped through torch.distributed."""

    def __init__(self):
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        uid = torch.zeros(128, dtype=torch.uint8)
        if self.rank == 0:
            buf = (C.c_char * 128)()
            _lib.check(lib().dvf_comm_unique_id(buf))
            uid = torch.frombuffer(bytearray(buf.raw), dtype=torch.uint8).clone()
        on_gpu = dist.get_backend() == "nccl"
        t = uid.cuda() if on_gpu else uid
        dist.broadcast(t, 0)
        raw = bytes(t.cpu().numpy().tobytes())
        self.handle = C.c_void_p()
        _lib.check(lib().dvf_comm_init(raw, self.rank, self.world, C.byref(self.handle)))

    def close(self):
        if self.handle:
            lib().dvf_comm_destroy(self.handle)
            self.handle = None


def gather_rows(t: torch.Tensor, dst: int = 0):
    """Concatenate the row blocks of all ranks on `dst` (ragged row counts allowed)."""
    world, rank = dist.get_world_size(), dist.get_rank()
    n = torch.tensor([t.shape[0]], dtype=torch.int64, device=t.device)
    counts = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(counts, n)
    counts = [int(c.item()) for c in counts]
    # equal-size gather (gloo and nccl both want it): pad every block to the largest row count
    mx = max(counts) if counts else 0
    if mx == 0:
        return (t[:0].clone() if rank == dst else None), counts
    pad = t.contiguous()
    if t.shape[0] < mx:
        pad = torch.cat([pad, torch.zeros((mx - t.shape[0],) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)], 0)
    parts = [torch.empty_like(pad) for _ in range(world)] if rank == dst else None
    dist.gather(pad, parts, dst=dst)
    if rank == dst:
        parts = [p[:c] for p, c in zip(parts, counts)]
    return (torch.cat(parts, 0) if rank == dst else None), counts


class RowServer:
    """Rows of the row-sharded matrix `Xv` by GLOBAL row number, asked for by rank 0 (a collective: every rank calls
    fetch() the same number of times; only rank 0 passes indices and gets rows back, in the order it asked)."""

    def __init__(self, Xv: torch.Tensor, counts):
        self.Xv, self.rank = Xv, dist.get_rank()
        self.lo = int(sum(counts[:self.rank]))
        self.hi = self.lo + int(counts[self.rank])

    def fetch(self, gidx=None):
        dev_ = self.Xv.device
        k = torch.tensor([gidx.numel() if self.rank == 0 else 0], dtype=torch.int64, device=dev_)
        dist.broadcast(k, 0)
        k = int(k.item())
        g = gidx.to(torch.int64).contiguous() if self.rank == 0 else torch.empty(k, dtype=torch.int64, device=dev_)
        if k:
            dist.broadcast(g, 0)
        sel = torch.nonzero((g >= self.lo) & (g < self.hi), as_tuple=False).reshape(-1)
        rows, _ = gather_rows(self.Xv[g[sel] - self.lo])
        pos, _ = gather_rows(sel)
        if self.rank != 0:
            return None
        out = torch.empty((k, self.Xv.shape[1]), dtype=self.Xv.dtype, device=dev_)
        if k:
            out[pos] = rows
        return out


def sharded_prelude(Xv, Yv, M, seed=0, beta=None, velocity_based_sampling=True):
    """Control points and bandwidth of the whole (row-sharded) data set, identical to prelude.run on the
    concatenated rows (tests/test_dist_cpu.py).  The row order np.unique defines (scVectorField.py:403) needs a
    global sort, but only of the first coordinate: rank 0 gathers that column and the row speeds (16 bytes per
    cell instead of 16 D), sorts, asks the owners for the few rows whose first coordinate is tied, samples on the
    host as the reference does, and asks the owners for the M chosen rows.
    Returns (ctrl_t, idx (numpy), beta, n_total) on every rank."""
    rank = dist.get_rank()
    D = Xv.shape[1]
    import os
    import sys
    import time
    timing = os.environ.get("DVF_PRELUDE_TIMING") and rank == 0   # developer aid: phase times on rank 0
    marks = []

    def mark(name):
        if timing:
            if Xv.is_cuda:
                torch.cuda.synchronize()
            marks.append((name, time.perf_counter()))

    mark("start")
    key, counts = gather_rows(Xv[:, 0].contiguous())
    speed = gather_rows(prelude.row_norms(Yv))[0] if velocity_based_sampling else None
    mark("gather keys+speeds")
    n_total = int(sum(counts))
    srv = RowServer(Xv, counts)
    meta = torch.zeros(2, dtype=torch.float64, device=Xv.device)
    if rank == 0:
        first = prelude.unique_first(key, srv.fetch)
        mark("sort + ties")
        n_u = first.numel()
        m_eff = min(int(M), n_u)
        if velocity_based_sampling:
            sp = prelude.host_f64(speed[first])
            mark("speeds to host")
            idx = prelude.velocity_weighted_choice(sp, m_eff, seed)
            mark("host choice")
        else:
            idx = np.random.RandomState(seed=seed).permutation(n_u)[:m_eff]  # scVectorField.py:409-410
        idx_t = torch.as_tensor(idx, dtype=torch.int64, device=Xv.device)
        ctrl_t = srv.fetch(first[idx_t])
        mark("fetch ctrl rows")
        if beta is None:
            beta = 1.0 / prelude.bandwidth(ctrl_t) ** 2
        mark("bandwidth")
        meta[0], meta[1] = float(m_eff), float(beta)
        if timing:
            print("[sharded prelude] n=%d: " % n_total + ", ".join("%s %.1f ms" % (b[0], (b[1] - a[1]) * 1e3)
                                                                    for a, b in zip(marks[:-1], marks[1:])),
                  file=sys.stderr, flush=True)
    else:
        srv.fetch()
        srv.fetch()
    dist.broadcast(meta, 0)
    m_eff, beta = int(meta[0].item()), float(meta[1].item())
    if rank != 0:
        ctrl_t = torch.empty((m_eff, D), dtype=torch.float64, device=Xv.device)
        idx_t = torch.empty((m_eff,), dtype=torch.int64, device=Xv.device)
    dist.broadcast(ctrl_t, 0)
    dist.broadcast(idx_t, 0)
    return ctrl_t, idx_t.cpu().numpy(), beta, n_total


def fit_sharded(X_local, Y_local, M, comm: Comm, seed=0, beta=None, velocity_based_sampling=True, **em):
    """SparseVFC over row-sharded cells: sharded_prelude fixes the control points and beta on every rank, then
    the EM loop runs on the local rows with one allreduce per iteration.
    Returns (FitState of the local rows, ctrl, idx, beta, local valid index)."""
    Xd, Yd = dev(X_local), dev(Y_local)
    valid = prelude.valid_rows(Yd)
    all_valid = valid.numel() == Yd.shape[0]
    Xv = Xd if all_valid else Xd[valid]
    Yv = Yd if all_valid else Yd[valid]
    ctrl_t, idx, beta, n_total = sharded_prelude(Xv, Yv, M, seed=seed, beta=beta,
                                                 velocity_based_sampling=velocity_based_sampling)
    st = fit_em(Xv, Yv, ctrl_t, beta, comm=comm.handle, N_total=n_total, **em)
    return st, ctrl_t, idx, beta, valid


def integrate_sharded(init_states, t, integration_direction, f, dst=0, **kwargs):
    """Trajectory integration over ranks ("replicas only": initial states are independent, there is
    no exchange step).  Every rank integrates its contiguous slice of `init_states` with integrate_vf_ivp on its own
    GPU; rank `dst` receives the lists of all ranks in the original order (other ranks get their own part back).
    `sampling="uniform_indices"` shares one time grid between all trajectories (prediction/utils.py:198-257) and is
    therefore not sharded."""
    from .integrate import integrate_vf_ivp

    if kwargs.get("sampling", "arc_length") == "uniform_indices":
        raise NotImplementedError("uniform_indices couples all trajectories; integrate them on one rank")
    world, rank = dist.get_world_size(), dist.get_rank()
    init_states = np.atleast_2d(np.asarray(init_states))
    lo, hi = shard_bounds(len(init_states), world, rank)
    tt, Y = integrate_vf_ivp(init_states[lo:hi], t, integration_direction, f, **kwargs) if hi > lo else ([], [])
    parts = [None] * world if rank == dst else None
    dist.gather_object((tt, Y), parts, dst=dst)
    if rank != dst:
        return tt, Y
    return [x for p in parts for x in p[0]], [y for p in parts for y in p[1]]


# ------------------------------------------------------------------------------------------------------
# The sharded fit behind the drop-in names (SURVEY.md section 5 "one env var / kwarg to select the devices")
# ------------------------------------------------------------------------------------------------------
_COMM = None


def default_comm() -> Comm:
    """One NCCL communicator per process over the default torch.distributed group, created on first use."""
    global _COMM
    if _COMM is None or not _COMM.handle:
        _COMM = Comm()
    return _COMM


def close_default_comm():
    global _COMM
    if _COMM is not None:
        _COMM.close()
        _COMM = None


def active() -> bool:
    """True when this process is one rank of a multi-rank torch.distributed job."""
    return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1


def SparseVFC_sharded(X, Y, Grid, M=100, a=5, beta=None, ecr=1e-5, gamma=0.9, lambda_=3, minP=1e-5, MaxIter=500,
                      theta=0.75, velocity_based_sampling=True, seed=0, lstsq_method="drouin", comm=None, gather=True,
                      dst=0):
    """SparseVFC (scVectorField.py:313-542) over cells sharded across the ranks of a torchrun job: `X`, `Y` are THIS
    rank's contiguous row block of the cell matrix (rank order = row order).  One joint fit: sharded_prelude fixes
    the control points of the whole data set, the EM loop allreduces the Gram/RHS block per iteration.

    gather=True: rank `dst` returns the VecFldDict of the whole data set -- the same keys, shapes and index
    conventions as a single-GPU SparseVFC of the concatenated rows (`X`, `Y`, `V`, `P`, `valid_ind`, `VFCIndex`
    gathered in row order).  Every other rank (and every rank with gather=False) returns the dict of its own rows:
    local `X`, `Y`, `V`, `P`, `valid_ind`, `VFCIndex` (local row numbers; `row_offset` gives the global number of
    its first row) next to the replicated `X_ctrl`, `ctrl_idx`, `beta`, `C`, `sigma2`, `iteration`, `tecr_traj`,
    `E_traj`, `grid`, `grid_V` -- enough for `vector_field_function` / Jacobian / divergence ... on its own cells."""
    from .device import host
    from .sparsevfc import solver_code
    from . import kernels

    comm = comm or default_comm()
    rank, world = dist.get_rank(), dist.get_world_size()
    solver = solver_code(lstsq_method)
    X_ori, Y_ori = np.array(X, copy=True), np.array(Y, copy=True)
    Xd, Yd = dev(X), dev(Y)
    st, ctrl_t, idx, beta, valid = fit_sharded(Xd, Yd, M, comm, seed=seed, beta=beta,
                                               velocity_based_sampling=velocity_based_sampling, a=a, gamma=gamma,
                                               lambda_=lambda_, minP=minP, ecr=ecr, theta=theta, MaxIter=MaxIter,
                                               solver=solver)
    if st.n_iter == 0:
        raise Exception(
            "please check your input parameters, "
            f"tecr: 1, ecr {ecr} and sigma2 {st.sigma2},"
            f"tecr must larger than ecr and sigma2 must larger than 1e-8"
        )
    i = st.n_iter
    n_rows = torch.tensor([Xd.shape[0]], dtype=torch.int64, device=Xd.device)
    all_rows = [torch.zeros_like(n_rows) for _ in range(world)]
    dist.all_gather(all_rows, n_rows)
    all_rows = [int(c.item()) for c in all_rows]
    offset = int(sum(all_rows[:rank]))
    ctrl = ctrl_t.cpu().numpy()
    shared = {
        "X_ctrl": ctrl, "ctrl_idx": idx, "beta": beta, "C": host(st.C), "sigma2": st.sigma2, "grid": Grid,
        "grid_V": None, "iteration": i - 1, "tecr_traj": host(st.tecr_traj)[:i], "E_traj": host(st.E_traj)[:i],
    }
    if Grid is not None:
        shared["grid_V"] = host(kernels.eval_field(Grid, ctrl_t, beta, st.C))
    if gather:
        Xg, _ = gather_rows(Xd, dst)
        Yg, _ = gather_rows(Yd, dst)
        Vg, _ = gather_rows(st.V, dst)
        Pg, _ = gather_rows(st.P.reshape(-1, 1), dst)
        vg, _ = gather_rows(valid + offset, dst)
        if rank == dst:
            P = host(Pg)
            return dict(shared, X=host(Xg), valid_ind=host(vg), Y=host(Yg), V=host(Vg), P=P,
                        VFCIndex=np.where(P > theta)[0])
    P = host(st.P).reshape(-1, 1)
    return dict(shared, X=X_ori, valid_ind=host(valid), Y=Y_ori, V=host(st.V), P=P, VFCIndex=np.where(P > theta)[0],
                row_offset=offset)
