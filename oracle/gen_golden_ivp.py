"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/ivp/ivp_cases.npz by EXECUTING THE UNMODIFIED REFERENCE
`integrate_vf_ivp` (prediction/utils.py:62-259) on small RKHS fields.

Run in the build container (needs /root/reference):   python -m oracle.gen_golden_ivp

Fields:  'sink'  -- SparseVFC fit (reference SvcVectorField.train) of V = -(X - x*) + small swirl, so trajectories
                    spiral into a fixed point and the terminal "at rest" event (:107) fires for long time spans;
         'rand'  -- random control points / coefficients in 6-D: trajectories leave the data and come to rest where
                    the kernel has decayed, or reach the end time first.
For every field: directions forward / backward / both x sampling arc_length / logspace / uniform_indices
(+ average False for uniform_indices), with interpolation_num 40.  The raw solver output of the reference's
solve_ivp call (t, y, status, sol at fixed fractions of the span) is stored too, for the kernel-level tests.
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle import ref_loader  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "ivp")
INTERP = 40
SOL_FRACTIONS = np.linspace(0.0, 1.0, 23)


def fields():
    ref_loader.install()
    sc, vfu, _ = ref_loader.hot_path()
    rng = np.random.default_rng(20240801)
    out = {}
    # sink: reference fit
    X = rng.normal(size=(600, 3)) * 1.5
    xs = np.array([0.4, -0.3, 0.2])
    R = np.array([[0, -1, 0], [1, 0, 0], [0, 0, 0.0]])
    V = -(X - xs) + 0.6 * (X - xs) @ R.T
    vf = sc.SvcVectorField(X, V, None, M=40, MaxIter=30, normalize=False).train()
    out["sink"] = dict(X_ctrl=np.asarray(vf["X_ctrl"]), C=np.asarray(vf["C"]), beta=float(vf["beta"]),
                       init=X[:5].copy(), t_end=np.array([3.0, 40.0]))
    # rand
    D, M = 6, 50
    out["rand"] = dict(X_ctrl=rng.normal(size=(M, D)), C=rng.normal(size=(M, D)) * 0.5, beta=0.35,
                       init=rng.normal(size=(4, D)) * 0.7, t_end=np.array([2.0, 60.0]))
    return out, vfu


def main():
    from scipy.integrate import solve_ivp

    os.makedirs(OUT, exist_ok=True)
    fs, vfu = fields()
    pu = ref_loader.load("dynamo.prediction.utils")
    z = {"interp": np.array(INTERP), "sol_fractions": SOL_FRACTIONS}
    for name, fd in fs.items():
        vf = {"X_ctrl": fd["X_ctrl"], "C": fd["C"], "beta": fd["beta"]}
        f = lambda x, vf=vf: vfu.vector_field_function(x, vf)  # noqa: E731
        for k in ("X_ctrl", "C", "init", "t_end"):
            z[f"{name}/{k}"] = fd[k]
        z[f"{name}/beta"] = np.array(fd["beta"])
        for te in fd["t_end"]:
            t = np.linspace(0, te, 7)
            # raw solver output, the call of :115-135
            for sign, tag in ((1.0, "fw"), (-1.0, "bw")):
                for i, y0 in enumerate(fd["init"]):
                    ev = lambda t_, x: np.all(abs(f(x)) < 1e-5) - 1  # noqa: E731
                    ev.terminal = True
                    r = solve_ivp(lambda t_, x: f(x), [sign * t[0], sign * t[-1]], y0, events=ev,
                                  max_step=np.abs(t[-1] - t[0]) / INTERP, dense_output=True)
                    key = f"{name}/T{te:g}/raw_{tag}/{i}"
                    z[key + "/t"], z[key + "/y"], z[key + "/status"] = r.t, r.y, np.array(r.status)
                    z[key + "/sol"] = r.sol(r.t[0] + SOL_FRACTIONS * (r.t[-1] - r.t[0]))
            for direction in ("forward", "backward", "both"):
                for sampling, average in (("arc_length", True), ("logspace", True), ("uniform_indices", True),
                                          ("uniform_indices", False)):
                    tt, Y = pu.integrate_vf_ivp(fd["init"], t, direction, f, interpolation_num=INTERP,
                                                average=average, sampling=sampling, disable=True)
                    key = f"{name}/T{te:g}/{direction}/{sampling}/avg{int(average)}"
                    z[key + "/n"] = np.array(len(Y))
                    for i in range(len(Y)):
                        z[f"{key}/t{i}"], z[f"{key}/Y{i}"] = np.asarray(tt[i]), np.asarray(Y[i])
        # prediction/fate.py:166-265 (_fate) on top of it: time sequence from getTseq, mean state per time point
        # (the loader's bare `dynamo.vectorfield` package skips the heavy __init__, so the one name fate.py imports
        # from it is bound by hand to the function that __init__ would have exported)
        sys.modules["dynamo.vectorfield"].vector_field_function = vfu.vector_field_function
        fate = ref_loader.load("dynamo.prediction.fate")
        for direction in ("forward", "both"):
            for sampling, average in (("arc_length", True), ("logspace", True), ("logspace", False)):
                tt, Y = fate._fate(f, fd["init"], t_end=float(fd["t_end"][0]), step_size=0.5, direction=direction,
                                   interpolation_num=INTERP, average=average, sampling=sampling)
                key = f"{name}/fate/{direction}/{sampling}/avg{int(average)}"
                z[key + "/n"] = np.array(len(Y))
                for i in range(len(Y)):
                    z[f"{key}/t{i}"], z[f"{key}/Y{i}"] = np.asarray(tt[i]), np.asarray(Y[i])
    path = os.path.join(OUT, "ivp_cases.npz")
    np.savez_compressed(path, **z)
    print("wrote", path, os.path.getsize(path), "bytes;", len(z), "arrays")
    st = {k: int(z[k]) for k in z if k.endswith("/status")}
    print("statuses:", {v: sum(1 for x in st.values() if x == v) for v in set(st.values())})
    print("steps:", sorted({len(z[k]) for k in z if k.endswith("/t") and "/raw_" in k}))


if __name__ == "__main__":
    main()
