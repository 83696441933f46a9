"""Loading of the committed golden vectors (tests/golden/*.npz, written by oracle/gen_golden.py from the
executed reference).  'lean' cases store a generator recipe + checksums instead of the inputs."""
import glob
import os

import numpy as np

from oracle import vfc_oracle

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(HERE, "golden", "*.npz")))


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


def prefix_len(a, b, tol=1e-8):
    """Number of leading entries of a that agree with b to `tol` (relative)."""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    n = 0
    for x, y in zip(a, b):
        if not abs(x - y) <= tol * abs(y):
            break
        n += 1
    return n


class Golden:
    def __init__(self, name):
        self.name = name
        self.z = np.load(os.path.join(HERE, "golden", name + ".npz"))
        self.lean = "gen" in self.z.files
        if self.lean:
            gen = self.z["gen"]
            if len(gen) == 5:     # bench.py's generator (manifold seed): the cases at the benchmark's own configurations
                from dynamo_release_b200.datasets import hematopoiesis_like

                n, d, seed, vscale, mseed = gen
                X, V = hematopoiesis_like(int(n), int(d), seed=int(seed), velocity_scale=float(vscale),
                                          manifold_seed=int(mseed))
            else:
                n, d, seed, vscale = gen
                X, V = vfc_oracle.hematopoiesis_like(int(n), int(d), seed=int(seed))
                V = V * vscale
            cs = np.array([X.sum(), V.sum(), np.abs(X).sum(), np.abs(V).sum()])
            assert np.allclose(cs, self.z["checksum"], rtol=1e-12), "regenerated inputs differ from the golden recipe"
            self.X, self.V = X, V
        else:
            self.X, self.V = self.z["X"], self.z["V"]
        self.params = {k: (int(v) if k in ("M", "MaxIter") else float(v))
                       for k, v in zip(self.z["params_keys"], self.z["params_vals"])}
        self.stride = 37 if self.lean else 1

    def __getitem__(self, k):
        return self.z[k]

    def vf(self):
        return {"X_ctrl": self.z["final_X_ctrl"], "beta": float(self.z["final_beta"]), "C": self.z["final_C"]}

    def kept_iters(self):
        return [int(k) for k in self.z["trace_kept_iters"]]

    def lhs_check(self, it, lhs):
        """Compare a full lhs against what the golden file holds for iteration `it` (the whole matrix, or -- big M --
        evenly spaced rows plus the column sums).  Returns the worst relative deviation."""
        if "trace_lhs_%d" % it in self.z.files:
            return rel(lhs, self.z["trace_lhs_%d" % it])
        rows = self.z["trace_lhs_rows_idx"]
        return max(rel(lhs[rows], self.z["trace_lhs_rows_%d" % it]), rel(lhs.sum(0), self.z["trace_lhs_colsum_%d" % it]))

    def floor_prefix(self, tol=1e-8):
        """How many leading EM iterations of the reference's own refit (cells permuted) agree with its first fit to
        `tol` in E -- the length of trajectory the reference itself can reproduce."""
        return prefix_len(self.z["floor_E_traj"], self.z["final_E_traj"], tol)

    def floor(self):
        c, v = float(self.z["floor_relC"]), float(self.z["floor_relV"])
        return (c if np.isfinite(c) else 0.0), (v if np.isfinite(v) else 0.0)
