"""TEST INFRASTRUCTURE ONLY -- golden vectors for the least-action-path functional from the EXECUTED reference
(vectorfield/least_action_path.py:7-120 and the same functions of prediction/least_action_path.py:260-420).

    python -m oracle.gen_golden_lap        (build container; needs /root/reference)  ->  tests/golden/post/lap.npz
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import ref_loader  # noqa: E402


def main():
    from golden_util import Golden

    ref_loader.install()
    utils = ref_loader.load("dynamo.vectorfield.utils")
    lap = ref_loader.load("dynamo.vectorfield.least_action_path")
    out = {}
    rng = np.random.default_rng(77)
    for tag, name in (("d30", "hema_n2000_d30_m100"), ("d12", "ragged_n1200_d12_m64")):
        g = Golden(name)
        vf = g.vf()
        f = lambda x, vf=vf: utils.vector_field_function(x, vf)          # noqa: E731
        jac = lambda x, vf=vf: utils.Jacobian_rkhs_gaussian(x, vf)       # noqa: E731
        cells = g["calc_cells"]
        a, b = g.X[cells[0]], g.X[cells[7]]
        P = 24
        path = np.linspace(0, 1, P)[:, None] * (b - a)[None, :] + a[None, :] + 0.05 * rng.standard_normal((P, len(a)))
        out[tag + "/path"] = path
        for D_, dt in ((1, 1), (0.5, 0.3)):
            out["%s/action_D%g_dt%g" % (tag, D_, dt)] = np.array(lap.action(path, f, D=D_, dt=dt))
            out["%s/grad_D%g_dt%g" % (tag, D_, dt)] = lap.action_grad(path, f, jac, D=D_, dt=dt)
        if tag == "d12":
            # vectorfield/least_action_path.py:115 hands scipy a 2-D x0, which scipy >= 1.11 refuses ("'x0' must only have
            # one dimension"); the minimisation is therefore driven as prediction/least_action_path.py:388 (lap_T) does it,
            # with the flattened interior points, around the reference's own action_aux / action_grad_aux
            from scipy.optimize import minimize
            n_points, dim = 8, len(a)
            path_0 = (np.tile(a, (n_points + 1, 1))
                      + (np.linspace(0, 1, n_points + 1, endpoint=True) * np.tile(b - a, (n_points + 1, 1)).T).T)
            res = minimize(lambda x: lap.action_aux(x, f, dim, start=path_0[0], end=path_0[-1], D=1),
                           path_0[1:-1].flatten(),
                           jac=lambda x: lap.action_grad_aux(x, f, jac, dim, start=path_0[0], end=path_0[-1], D=1))
            sol = lap.reshape_path(res["x"], dim, start=path_0[0], end=path_0[-1])
            out[tag + "/lap_start"], out[tag + "/lap_end"] = a, b
            out[tag + "/lap_path"], out[tag + "/lap_fun"] = sol, np.array(res["fun"])
            out[tag + "/lap_nit"] = np.array(res["nit"])
    dst = os.path.join(ROOT, "tests", "golden", "post", "lap.npz")
    np.savez_compressed(dst, **out)
    print("wrote", dst, {k: np.shape(v) for k, v in out.items()})


if __name__ == "__main__":
    main()
