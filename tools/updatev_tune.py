"""Developer tool: time dvf_update_v (V = U C, r2, sum P r2) alone."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dynamo_release_b200 import kernels
torch.manual_seed(0)
for N, D, M in ((1000000, 30, 500), (1000000, 50, 1000), (125000, 30, 500)):
    W = kernels.row_width(M, D)
    UY = torch.rand((N, W), dtype=torch.float64, device="cuda")
    C = torch.randn((M, D), dtype=torch.float64, device="cuda")
    P = torch.rand((N + 2,), dtype=torch.float64, device="cuda")
    f = lambda: kernels.update_v(UY, C, P, 1e-5, D, M)
    f(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(40):
        f()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 40
    print("update_v N=%d D=%d M=%d: %.3f ms (incl. pad + partial sums + allocations), %.0f GB/s of U" % (N, D, M, ms, N * ((M + 31) // 32 * 32) * 8 / ms / 1e6), flush=True)
