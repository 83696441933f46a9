"""Developer tool: time the fused divergence / acceleration / curvature kernel (dvf_div_acc_curv)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dynamo_release_b200 import kernels

torch.manual_seed(0)
for n, D, M in ((1000000, 30, 500), (1000000, 50, 1000), (1000000, 40, 500), (1000000, 64, 500), (1000000, 33, 200)):
    x = torch.randn((n, D), dtype=torch.float64, device="cuda")
    Xc = x[:M].clone() + 0.01
    C = torch.randn((M, D), dtype=torch.float64, device="cuda")
    f = lambda: kernels.div_acc_curv(x, Xc, 0.02, C)
    f(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        f()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    print("dac n=%d D=%d M=%d: %.2f ms  %.2f TF (12MD+M+8D per cell)" % (n, D, M, ms, n * (12.0 * M * D + M + 8 * D) / ms / 1e9), flush=True)
