"""Helpers shared by the IVP tests: the golden cases of the executed reference integrate_vf_ivp
(tests/golden/ivp/ivp_cases.npz, oracle/gen_golden_ivp.py) and an oracle-backed stand-in for the device solver."""
import os

import numpy as np

from oracle import ivp_oracle, vfc_oracle

HERE = os.path.dirname(os.path.abspath(__file__))
FIELDS = ("sink", "rand")
DIRECTIONS = ("forward", "backward", "both")
SAMPLINGS = (("arc_length", True), ("logspace", True), ("uniform_indices", True), ("uniform_indices", False))


def load():
    return np.load(os.path.join(HERE, "golden", "ivp", "ivp_cases.npz"))


def field(z, name):
    return {"X_ctrl": z[f"{name}/X_ctrl"], "C": z[f"{name}/C"], "beta": float(z[f"{name}/beta"])}


def combos(z):
    for name in FIELDS:
        for te in z[f"{name}/t_end"]:
            for direction in DIRECTIONS:
                for sampling, average in SAMPLINGS:
                    yield name, float(te), direction, sampling, average


def oracle_batch(dense_cls, result_cls):
    """solve_ivp_batch with the NumPy oracle underneath (CPU tests of the host-side sampling logic only)."""

    def run(init_states, vf_dict, t_span, max_step, scale=1.0, **kw):
        out = []
        for y0 in np.atleast_2d(init_states):
            r = ivp_oracle.rk45_solve(lambda x: scale * vfc_oracle.field_eval(x, vf_dict), t_span[0], t_span[1], y0,
                                      max_step=max_step)
            out.append(result_cls(r["t"], r["y"], dense_cls(r["t"], r["y"], r["Q"], r["h"]), r["status"]))
        return out

    return run


FATE_CASES = [(d, s, a) for d in ("forward", "both") for s, a in (("arc_length", True), ("logspace", True),
                                                                  ("logspace", False))]


def check_fate(z, it, tol):
    """_fate (prediction/fate.py:166-265) for every golden case; returns the worst relative difference."""
    from golden_util import rel

    worst = 0.0
    for name in FIELDS:
        f = it.RkhsField(field(z, name))
        for direction, sampling, average in FATE_CASES:
            t, Y = it._fate(f, z[f"{name}/init"], t_end=float(z[f"{name}/t_end"][0]), step_size=0.5,
                            direction=direction, interpolation_num=int(z["interp"]), average=average,
                            sampling=sampling)
            key = f"{name}/fate/{direction}/{sampling}/avg{int(average)}"
            assert len(Y) == int(z[key + "/n"]), key
            for i in range(len(Y)):
                assert Y[i].shape == z[f"{key}/Y{i}"].shape and np.shape(t[i]) == z[f"{key}/t{i}"].shape, key
                e = max(rel(t[i], z[f"{key}/t{i}"]), rel(Y[i], z[f"{key}/Y{i}"]))
                assert e < tol, (key, i, e)
                worst = max(worst, e)
    return worst
