"""Developer tool: time the per-cell post-processing kernels (K9) and the fused ranking reductions.
Usage: python tools/post_tune.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dynamo_release_b200 import kernels


def timed(fn, reps=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


torch.manual_seed(0)
for D, n in ((30, 1000000), (50, 400000), (10, 1000000), (64, 200000)):
    J = torch.randn((D, D, n), dtype=torch.float64, device="cuda")
    ms = timed(lambda: kernels.jacobian_det(J))
    print("jacobian_det D=%d n=%d: %.3f ms  %.0f GB/s  %.2f M cells/s" % (D, n, ms, J.numel() * 8 / ms / 1e6, n / ms / 1e3), flush=True)
    del J
for D, n in ((30, 1000000), (50, 1000000)):
    U = torch.randn((n, D), dtype=torch.float64, device="cuda"); V = torch.randn((n, D), dtype=torch.float64, device="cuda")
    ms = timed(lambda: kernels.cell_angles(U, V))
    print("cell_angles D=%d n=%d: %.3f ms  %.0f GB/s" % (D, n, ms, 2 * U.numel() * 8 / ms / 1e6), flush=True)
    ms = timed(lambda: kernels.row_norms(V))
    print("row_norms   D=%d n=%d: %.3f ms  %.0f GB/s" % (D, n, ms, U.numel() * 8 / ms / 1e6), flush=True)
D, G, n = 50, 2000, 4096
J = torch.randn((D, D, n), dtype=torch.float64, device="cuda")
Q = torch.randn((G, D), dtype=torch.float64, device="cuda")
lab = torch.randint(0, 10, (n,), dtype=torch.int32, device="cuda")
flops = 2.0 * G * D * (D + G) * n
for op in ("sum", "switch", "max"):
    ms = timed(lambda: kernels.gene_transform_reduce(J, Q, Q, op=op, group=lab, n_groups=10), reps=2)
    print("gene reduce %-6s D=%d G=%d n=%d: %.2f ms  %.2f TF  %.0f cells/s" % (op, D, G, n, ms, flops / ms / 1e9, n / ms * 1e3), flush=True)
jm = kernels.gene_transform_reduce(J, Q, Q, op="max").to(torch.float32)
ms = timed(lambda: kernels.gene_transform_reduce(J, Q, Q, op="sum", jmax=jm, group=lab, n_groups=10), reps=2)
print("gene reduce sum+normalize: %.2f ms  %.0f cells/s" % (ms, n / ms * 1e3), flush=True)
