"""Throughput of K11 (trajectories/s and field evaluations/s) next to the oracle's per-trajectory time on the host.
    python tools/ivp_tune.py [n_traj] [D] [M] [interp]"""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from dynamo_release_b200 import integrate as it  # noqa: E402
from oracle import ivp_oracle as io, vfc_oracle as vo  # noqa: E402

n, D, M, interp = (int(a) for a in (sys.argv[1:5] + ["4096", "30", "500", "250"][len(sys.argv) - 1:]))
rng = np.random.default_rng(0)
vf = {"X_ctrl": rng.normal(size=(M, D)), "C": rng.normal(size=(M, D)) * 0.3, "beta": 0.5 / D}
y0 = rng.normal(size=(n, D))
T = 3.0
it.PROFILE = True
for rep in range(3):
    it.KERNEL_MS = 0.0
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    res = it.solve_ivp_batch(y0, vf, (0.0, T), T / interp)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
steps = sum(len(r.t) - 1 for r in res)
print(f"K11 n={n} D={D} M={M} interp={interp}: kernel {it.KERNEL_MS:.1f} ms = {n/it.KERNEL_MS*1e3:.0f} traj/s, "
      f"{6*steps/it.KERNEL_MS/1e3:.2f} M field evaluations/s, {6*steps*(5*M*D+M)/it.KERNEL_MS/1e9:.2f} TFLOP/s algorithmic; "
      f"solve_ivp_batch {dt*1e3:.1f} ms (times / states to the host, result objects) = {n/dt:.0f} traj/s; "
      f"{steps} accepted steps")
f_dev = it.RkhsField(vf)
for sampling in ("arc_length", "logspace"):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    tt, Y = it.integrate_vf_ivp(y0, np.linspace(0, T, 7), "forward", f_dev, interpolation_num=interp, sampling=sampling,
                                disable=True)
    torch.cuda.synchronize()
    de = time.perf_counter() - t0
    print(f"  integrate_vf_ivp(sampling={sampling}): {de*1e3:.1f} ms = {n/de:.0f} traj/s")
# kernel alone
ts = time.perf_counter()
f = lambda x: vo.field_eval(x, vf)  # noqa: E731
k = 8
for i in range(k):
    io.rk45_solve(f, 0.0, T, y0[i], max_step=T / interp)
dc = (time.perf_counter() - ts) / k
print(f"oracle solve (NumPy, 1 thread): {dc*1e3:.1f} ms per trajectory = {1/dc:.1f} traj/s; solve_ivp_batch ratio {n/dt*dc:.0f}x")
