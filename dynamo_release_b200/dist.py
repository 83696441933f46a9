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
    pad = t.contiguous()
    if t.shape[0] < mx:
        pad = torch.cat([pad, torch.zeros((mx - t.shape[0],) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)], 0)
    parts = [torch.empty_like(pad) for _ in range(world)] if rank == dst else None
    dist.gather(pad, parts, dst=dst)
    if rank == dst:
        parts = [p[:c] for p, c in zip(parts, counts)]
    return (torch.cat(parts, 0) if rank == dst else None), counts


def fit_sharded(X_local, Y_local, M, comm: Comm, seed=0, beta=None, velocity_based_sampling=True, **em):
    """SparseVFC over row-sharded cells.  The prelude needs the global row order (np.unique semantics,
    scVectorField.py:403-411): valid rows are gathered on rank 0, which samples the control points and
    broadcasts them with beta.  Returns (FitState of the local rows, ctrl, idx, beta, local valid index)."""
    Xd, Yd = dev(X_local), dev(Y_local)
    valid = prelude.valid_rows(Yd)
    Xv, Yv = Xd[valid], Yd[valid]
    Xg, counts = gather_rows(Xv)
    Yg, _ = gather_rows(Yv)
    n_total = int(sum(counts))
    D = Xd.shape[1]
    meta = torch.zeros(2, dtype=torch.float64, device=Xd.device)
    if comm.rank == 0:
        _, _, _, ctrl, idx, beta = prelude.run(Xg, Yg, M, seed=seed, beta=beta,
                                               velocity_based_sampling=velocity_based_sampling)
        meta[0], meta[1] = float(ctrl.shape[0]), beta
        ctrl_t = dev(ctrl)
        idx_t = torch.as_tensor(idx, dtype=torch.int64, device=Xd.device)
    dist.broadcast(meta, 0)
    m_eff, beta = int(meta[0].item()), float(meta[1].item())
    if comm.rank != 0:
        ctrl_t = torch.empty((m_eff, D), dtype=torch.float64, device=Xd.device)
        idx_t = torch.empty((m_eff,), dtype=torch.int64, device=Xd.device)
    dist.broadcast(ctrl_t, 0)
    dist.broadcast(idx_t, 0)
    st = fit_em(Xv, Yv, ctrl_t, beta, comm=comm.handle, N_total=n_total, **em)
    return st, ctrl_t, idx_t.cpu().numpy(), beta, valid
