#!/usr/bin/env python
"""bench.py -- cells/s for SparseVFC fit + per-cell Jacobian + divergence / acceleration / curvature.

Workload (BASELINE.json configs[2], the one the metric is quoted on; fits one GPU): synthetic hematopoiesis-shaped
1M cells x 30 PCs per GPU, M = 500 control points, MaxIter = 30.  N > 1: weak scaling, every rank holds 1M cells,
ONE joint fit over N x 1M cells (NCCL allreduce of the Gram/RHS block per EM iteration), evaluation per rank.

  value : whole-job cells/s with the inputs resident in HBM (CUDA events, max over ranks)
  e2e   : the same job through the public numpy API (pinned host inputs -> numpy outputs, copies timed)
  roofline : the dominant kernel (weighted Gram, FP64 tensor cores), CUDA-event timed inside the fit
  cpu_baseline / --impl reference : the reference's algorithm (oracle port, NumPy/BLAS, all host threads) on a
      bounded sample of the same workload.
One JSON line on stdout (rank 0).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GRAM_DRAM_BYTES_NCU = 4.85e9    # k_gram_parts: 4.84 GB read + 7 MB written per launch (ncu --set full, profiles/ncu_k_gram_parts_r01_details.txt)
FP64_PEAK_TFLOPS = 37.1   # DMMA m8n8k4 chain measured on this pool's B200 (tools/fp64_peak.cu, profiles/fp64_peak_r01.json)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cells", type=int, default=1_000_000, help="cells per GPU")
    ap.add_argument("--pcs", type=int, default=30)
    ap.add_argument("--ctrl", type=int, default=500)
    ap.add_argument("--maxiter", type=int, default=30)
    ap.add_argument("--velocity-scale", type=float, default=0.05)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def workload_name(a):
    return ("synthetic hematopoiesis-shaped %d cells x %d PCs per GPU, M=%d ctrl pts, MaxIter=%d: SparseVFC fit + "
            "jacobian + divergence + acceleration + curvature" % (a.cells, a.pcs, a.ctrl, a.maxiter))


# ----------------------------------------------------------------------------------------------------------
# CPU baseline: the reference's algorithm (oracle/vfc_oracle.py, pinned against the executed reference by
# tests/test_oracle_golden.py) on a bounded sample, all host threads available to BLAS.
# ----------------------------------------------------------------------------------------------------------
def cpu_reference_rate(a, n_iter_full=None, n_fit=50_000, n_eval=1_500):
    from oracle import vfc_oracle as o
    from dynamo_release_b200.datasets import hematopoiesis_like

    X, V = hematopoiesis_like(n_fit, a.pcs, seed=1, velocity_scale=a.velocity_scale, manifold_seed=7)
    t0 = time.perf_counter()
    vf = o.train(X, V, M=a.ctrl, MaxIter=a.maxiter)
    t_fit = time.perf_counter() - t0
    iters = vf["iteration"] + 1
    xs = X[:n_eval]
    t0 = time.perf_counter()
    o.jacobian(xs, vf)
    t_jac = time.perf_counter() - t0
    t0 = time.perf_counter()
    o.divergence(xs, vf)
    t_div = time.perf_counter() - t0
    t0 = time.perf_counter()
    o.acceleration(xs, vf)
    t_acc = time.perf_counter() - t0
    t0 = time.perf_counter()
    o.curvature(xs, vf)
    t_cur = time.perf_counter() - t0
    # per-cell cost, linear in the number of cells (per-cell loops / N-row GEMMs); the fit is rescaled to the
    # iteration count of the full-size run when it is known
    fit_scale = (n_iter_full / iters) if n_iter_full else 1.0
    per_cell = t_fit * fit_scale / n_fit + (t_jac + t_div + t_acc + t_cur) / n_eval
    sample = ("fit on %d cells (%d EM iterations, %.2f s%s) + jacobian/divergence/acceleration/curvature on %d "
              "cells (%.2f/%.2f/%.2f/%.2f s); per-cell costs added, linear extrapolation in cells"
              % (n_fit, iters, t_fit, (", rescaled to %d iterations" % n_iter_full) if n_iter_full else "", n_eval,
                 t_jac, t_div, t_acc, t_cur))
    return 1.0 / per_cell, sample, t_fit + t_jac + t_div + t_acc + t_cur


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    vals, sample, ms = [], "", []
    for i in range(a.warmup + a.steps):
        n_fit, n_eval = (10_000, 300) if i < a.warmup else (30_000, 1_000)
        t0 = time.perf_counter()
        rate, sample, _ = cpu_reference_rate(a, None, n_fit, n_eval)
        if i >= a.warmup:
            vals.append(rate)
            ms.append((time.perf_counter() - t0) * 1e3)
    v = float(np.mean(vals))
    line = {
        "impl": "reference", "metric": "cells/sec for SparseVFC fit + per-cell Jacobian/divergence/acceleration/curvature",
        "value": v, "unit": "cells/s", "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": float(np.mean(ms)), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(a), "note": "reference algorithm (NumPy/SciPy oracle port of dynamo's "
                   "CPU path) on the host cores; each step is a bounded sample of the workload"},
        "cpu_baseline": {"value": v, "unit": "cells/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "cells/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.rows.append([c.strip() for c in ln.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], None, set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
                pw.append(float(r[2]))
                for nm, val in zip(names, r[3:7]):
                    if val.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        busy = [s for s, p in zip(sm, pw) if p > 250] or sm
        return {"sm_mhz": float(np.median(busy)) if busy else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(pw) if pw else None}


# ----------------------------------------------------------------------------------------------------------
def run_ours(a):
    import torch
    import torch.distributed as dist

    import dynamo_release_b200 as dvf
    from dynamo_release_b200 import _lib, kernels, prelude
    from dynamo_release_b200.datasets import hematopoiesis_like
    from dynamo_release_b200.device import pinned_like
    from dynamo_release_b200.sparsevfc import fit_em

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    comm = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        from dynamo_release_b200.dist import Comm, fit_sharded
        comm = Comm()
    lib = _lib.load()
    N, D, M = a.cells, a.pcs, a.ctrl
    Xh, Vh = hematopoiesis_like(N, D, seed=100 + rank, velocity_scale=a.velocity_scale, manifold_seed=7)
    Xh, Vh = pinned_like(Xh), pinned_like(Vh)
    Xd, Vd = torch.from_numpy(Xh).cuda(), torch.from_numpy(Vh).cuda()
    em = dict(MaxIter=a.maxiter)
    J_buf = torch.empty((D, D, N), dtype=torch.float64, device="cuda")
    state = {}

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    stage_ev = []

    def mark():
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        return e

    def step_device():
        """whole job, inputs resident in HBM, results left in HBM"""
        ev = [mark()]
        if world > 1:
            st, ctrl, _, beta, _ = fit_sharded(Xd, Vd, M, comm, **em)
            ev.append(ev[0])
        else:
            _, Xv, Yv, ctrl, _, beta = prelude.run(Xd, Vd, M, seed=0)
            ev.append(mark())
            st = fit_em(Xv, Yv, ctrl, beta, workspace=state.get("ws"), **em)
            state["ws"] = st.workspace
        ev.append(mark())
        kernels.jacobian(Xd, ctrl, beta, st.C, out=J_buf)
        ev.append(mark())
        r = kernels.div_acc_curv(Xd, ctrl, beta, st.C, formula=2)
        ev.append(mark())
        stage_ev.append(ev)
        state.update(n_iter=st.n_iter, sigma2=st.sigma2, n_inlier=st.n_inlier, r=r)

    def step_e2e():
        """public numpy API: pinned host arrays in, numpy arrays out (single GPU path of the drop-in)"""
        t = [time.perf_counter()]
        vfc = dvf.SvcVectorField(Xh, Vh, None, normalize=False, M=M, **em)
        vf = vfc.train()
        t.append(time.perf_counter())
        J = dvf.Jacobian_rkhs_gaussian(Xh, vf)
        t.append(time.perf_counter())
        div = vfc.compute_divergence()
        t.append(time.perf_counter())
        acc, acc_mat = vfc.compute_acceleration()
        t.append(time.perf_counter())
        curv, curv_mat = vfc.compute_curvature()
        t.append(time.perf_counter())
        state["e2e_stages_ms"] = {k: (t[i + 1] - t[i]) * 1e3 for i, k in enumerate(
            ["train", "jacobian", "divergence", "acceleration", "curvature"])}
        h2d = 6 * Xh.nbytes                                 # X, V for the fit + X for each of the 4 evaluations
        d2h = (vf["V"].nbytes + vf["P"].nbytes + vf["C"].nbytes + J.nbytes + div.nbytes + acc.nbytes +
               acc_mat.nbytes + curv.nbytes + curv_mat.nbytes)
        return h2d, d2h

    # ---- device-resident timing ----
    for _ in range(a.warmup):
        step_device()
    lib.dvf_profile_enable(1)
    stage_ev.clear()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        step_device()
    e1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms = e0.elapsed_time(e1) / a.steps
    gram_ms, gram_n, launches = C.c_double(0), C.c_int(0), C.c_longlong(0)
    lib.dvf_profile_get(C.byref(gram_ms), C.byref(gram_n), C.byref(launches))
    lib.dvf_profile_enable(0)
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    value = world * N / (ms * 1e-3)

    # ---- end-to-end through the numpy API (rank-local job; with N ranks, N independent API calls) ----
    e2e = None
    if not a.no_e2e:
        for _ in range(max(1, min(a.warmup, 2))):
            h2d, d2h = step_e2e()
        barrier()
        t0 = time.perf_counter()
        for _ in range(a.steps):
            h2d, d2h = step_e2e()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / a.steps
        t = torch.tensor([dt], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e = {"value": world * N / float(t.item()), "unit": "cells/s", "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h), "ms_per_step": float(t.item()) * 1e3,
               "stages_ms_last_step": state.get("e2e_stages_ms"),
               "note": "numpy API, pinned host inputs, numpy outputs incl. the %.1f GB Jacobian; N>1: one API job "
                       "per rank" % (J_buf.numel() * 8 / 1e9)}

    if rank == 0:
        n_iter = state["n_iter"]
        n_tot = world * N
        # algorithmic flops of one Gram launch: symmetric Gram M^2 + M, RHS 2 M D per cell
        gram_flops = float(N) * (M * M + M + 2.0 * M * D)
        per_launch_ms = gram_ms.value / max(1, a.steps * n_iter)   # launches after the stop flag return at once
        achieved = gram_flops / (per_launch_ms * 1e-3) / 1e12 if per_launch_ms > 0 else 0.0
        line = {
            "metric": "cells/sec for SparseVFC fit + per-cell Jacobian/divergence/acceleration/curvature",
            "value": value, "unit": "cells/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": workload_name(a), "cells_total": n_tot, "em_iterations": n_iter,
                       "sigma2": state["sigma2"], "inlier_frac": state["n_inlier"] / n_tot,
                       "l2": "inputs larger than L2 (UY %.1f GB, J %.1f GB per GPU)"
                             % (N * kernels.row_width(M, D) * 8 / 1e9, J_buf.numel() * 8 / 1e9),
                       "parallelism": "cells sharded over %d GPU(s); Gram/RHS allreduce per EM iteration" % world},
            "clocks": clocks,
            "stages_ms": {k: float(np.mean([ev[i].elapsed_time(ev[i + 1]) for ev in stage_ev]))
                          for i, k in enumerate(["prelude", "fit", "jacobian", "div_acc_curv"])},
            "e2e": e2e,
            "gpu_launches": int(launches.value),
            "roofline": {"bound": "tensor", "kernel": "k_gram_parts (k_gram for many tiles): weighted Gram + RHS, FP64 DMMA m8n8k4 fed by TMA, + its reduce kernel",
                         "achieved": achieved, "peak": FP64_PEAK_TFLOPS, "unit": "TFLOP/s",
                         "frac": achieved / FP64_PEAK_TFLOPS,
                         # dram__bytes_read.sum + dram__bytes_write.sum of one k_gram launch at the default workload from
                         # the committed ncu --set full capture (profiles/ncu_k_gram_r01_details.txt); null elsewhere
                         "traffic": GRAM_DRAM_BYTES_NCU if (N, D, M) == (1_000_000, 30, 500) else None,
                         "traffic_unit": "bytes per launch (compulsory: %d)" % (N * kernels.row_width(M, D) * 8),
                         "ms_per_launch": per_launch_ms, "launches_timed": a.steps * n_iter,
                         "share_of_step": gram_ms.value / a.steps / ms,
                         "peak_source": "FP64 DMMA peak measured on this pool (tools/fp64_peak.cu: 37.1 TFLOP/s; "
                                        "cuBLAS DGEMM 35.5); MEASURED_PEAKS.json has no FP64 entry",
                         "algorithmic_flops_per_launch": gram_flops,
                         "hbm_GBps_compulsory": N * kernels.row_width(M, D) * 8 / (per_launch_ms * 1e-3) / 1e9
                         if per_launch_ms > 0 else None},
        }
        if not a.no_cpu_baseline and world == 1:
            rate, sample, _ = cpu_reference_rate(a, n_iter)
            line["cpu_baseline"] = {"value": rate, "unit": "cells/s", "cores": os.cpu_count(), "kind": "port",
                                    "sample": sample}
        print(json.dumps(line), flush=True)
    if world > 1:
        comm.close()
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
