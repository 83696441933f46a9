"""Developer tool: which entries of the float32 gene-space Jacobian differ from the executed reference, and by how much."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from dynamo_release_b200 import kernels
from oracle import vfc_oracle as o
# the bench's own last slab: c4_inputs(N = 200000, seed 2), cells 198656 .. 199999, compared on its first 48 cells
rng = np.random.default_rng(2); D, G, N = 50, 2000, 200000
Q, _ = np.linalg.qr(rng.normal(size=(G, D))); Q = np.ascontiguousarray(Q); Jall = rng.standard_normal((D, D, N))
Jslab = np.ascontiguousarray(Jall[:, :, 198656:]); n = 48
J = np.ascontiguousarray(Jslab[:, :, :n])
want = o.gene_jacobian(J, Q, Q)
want64 = np.einsum("ei,ijc,rj->erc", Q, J, Q, optimize=True)
got = kernels.gene_transform(Jslab, Q, Q)[0][:, :, :n].cpu().numpy()
again = kernels.gene_transform(Jslab, Q, Q)[0][:, :, :n].cpu().numpy()
print("two launches bit-equal:", np.array_equal(got, again))
alone = kernels.gene_transform(J, Q, Q)[0].cpu().numpy()
print("slab of 1344 vs the 48 cells alone bit-equal:", np.array_equal(got, alone))
ulp = np.abs(got.astype(np.float64) - want.astype(np.float64)) / np.spacing(np.abs(want)).astype(np.float64)
print("differ", int((ulp > 0).sum()), "beyond 1 ulp", int((ulp > 1).sum()), "max", ulp.max())
for e, r, c in np.argwhere(ulp > 0)[:12]:
    print(" got %.9e ref32 %.9e f64 %.17e | cell max %.2e | ulp %.1f | got-f64 %.2e ref-f64 %.2e" % (
        got[e, r, c], want[e, r, c], want64[e, r, c], np.abs(want[:, :, c]).max(), ulp[e, r, c],
        float(got[e, r, c]) - want64[e, r, c], float(want[e, r, c]) - want64[e, r, c]))
