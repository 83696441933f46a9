"""Developer tool: time dvf_jacobian alone and compare the GEMM form with the direct per-cell form (DVF_JAC_DIRECT=1
in a subprocess) and with a float64 torch evaluation of the reference formula on a cell subset."""
import os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def one(N, M, D):
    import torch
    from dynamo_release_b200 import kernels
    torch.manual_seed(0)
    x = torch.randn((N, D), dtype=torch.float64, device="cuda") * 2 + 5.0      # off-centre on purpose
    Xc = x[torch.randperm(N, device="cuda")[:M]].clone() + 0.01
    C = torch.randn((M, D), dtype=torch.float64, device="cuda")
    beta = 0.05
    J = torch.empty((D, D, N), dtype=torch.float64, device="cuda")
    for _ in range(2):
        kernels.jacobian(x, Xc, beta, C, out=J)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 3
    e0.record()
    for _ in range(reps):
        kernels.jacobian(x, Xc, beta, C, out=J)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    # reference formula on a subset (cells spread over the range, incl. the last ones)
    idx = torch.unique(torch.cat([torch.arange(0, N, max(1, N // 2000), device="cuda"), torch.arange(max(0, N - 200), N, device="cuda")]))
    xs = x[idx]
    diff = xs[:, None, :] - Xc[None, :, :]                      # n x M x D
    k = torch.exp(-beta * (diff * diff).sum(-1))                # n x M
    Jr = -2 * beta * torch.einsum("nm,mi,nmj->ijn", k, C, diff)
    err = float((J[:, :, idx] - Jr).abs().max() / Jr.abs().max())
    flops = N * (2.0 * M * D * D + 4.0 * M * D + M + D * D)
    print("jacobian N=%d M=%d D=%d  %.3f ms  %.2f TF algorithmic  relerr=%.2e  direct=%s" % (
        N, M, D, ms, flops / ms / 1e9, err, os.environ.get("DVF_JAC_DIRECT", "0")), flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 3:
        one(int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]))
    else:
        for cfg in [(1000000, 500, 30), (999937, 500, 30), (100000, 100, 30), (5000, 100, 2), (200000, 200, 50), (300, 20, 3),
                    (20000, 37, 10), (50000, 64, 64), (1, 5, 3)]:
            for direct in ("0", "1"):
                subprocess.run([sys.executable, __file__] + [str(c) for c in cfg], env=dict(os.environ, DVF_JAC_DIRECT=direct))
