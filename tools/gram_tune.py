"""Developer tool: time the Gram kernel (dvf_gram) alone for different schedule knobs (env DVF_GRAM_*) and check
it against torch's FP64 matmul on a row subset.  Usage: python tools/gram_tune.py [N M D] | sweep"""
import os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def one(N, M, D, check=True):
    import torch
    from dynamo_release_b200 import kernels
    torch.manual_seed(0)
    W = kernels.row_width(M, D)
    UY = torch.rand((N, W), dtype=torch.float64, device="cuda")
    UY[:, M + D:] = 0
    P = torch.rand((N,), dtype=torch.float64, device="cuda")
    for _ in range(2):
        G = kernels.gram(UY, P, 1e-5, D, M)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    e0.record()
    for _ in range(reps):
        G = kernels.gram(UY, P, 1e-5, D, M)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    err = float("nan")
    nondet = 0
    G0 = G.clone()
    for _ in range(6):
        nondet += int(not torch.equal(kernels.gram(UY, P, 1e-5, D, M), G0))
    if check:
        Gr = torch.zeros((W, W), dtype=torch.float64, device="cuda")
        for i in range(0, N, 100000):
            u = UY[i:i + 100000]
            Gr += (u * P[i:i + 100000, None].clamp_min(1e-5)).T @ u
        K = M + D
        tri = torch.tril(torch.ones((K, K), dtype=torch.bool, device="cuda"))
        err = float(((G[:K, :K] - Gr[:K, :K]).abs() * tri).max() / Gr.abs().max())
    flops = N * (M * M + M + 2.0 * M * D)
    print("N=%d M=%d D=%d  %.3f ms (incl. clamp copy)  %.2f TF algorithmic  relerr=%.2e nondet=%d/6 env=%s" % (
        N, M, D, ms, flops / ms / 1e9, err, nondet, {k: v for k, v in os.environ.items() if k.startswith("DVF_GRAM")}), flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 3:
        one(int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), check=not os.environ.get("DVF_GRAM_NOCHECK"))
    else:
        big = (1000000, 500, 30)
        runs = [({"DVF_GRAM_DEBUG": "1"}, (37, 20, 3)), ({"DVF_GRAM_DEBUG": "1"}, (1000, 500, 30)),
                ({"DVF_GRAM_DEBUG": "1"}, big), ({}, (999983, 500, 30)), ({"DVF_GRAM_LAG": "100000"}, big),
                ({"DVF_GRAM_LAG": "0"}, big), ({"DVF_GRAM_LAG": "2"}, big),
                ({"DVF_GRAM_SB": "4"}, big), ({"DVF_GRAM_SB": "8"}, big), ({"DVF_GRAM_SB": "32"}, big),
                ({"DVF_GRAM_SB": "64", "DVF_GRAM_LAG": "0"}, big)]
        for v in range(7):
            runs.append(({"DVF_GRAM_ONLY": str(v), "DVF_GRAM_DEBUG": "1", "DVF_GRAM_NOCHECK": "1", "DVF_GRAM_G": "1"},
                         (1000000, 500 + (32 if v in (2, 3) else 0), 30)))
        for G in (1, 9, 11, 12, 14):
            runs.append(({"DVF_GRAM_DEBUG": "1", "DVF_GRAM_G": str(G)}, big))
        runs.append(({"DVF_GRAM_DEBUG": "1"}, (1000000, 100, 30)))
        runs.append(({"DVF_GRAM_DEBUG": "1"}, (100000, 100, 30)))
        runs.append(({"DVF_GRAM_DEBUG": "1"}, (1000000, 1000, 50)))
        runs.append(({"DVF_GRAM_DEBUG": "1"}, (1000000, 200, 50)))
        for env, (N, M, D) in runs:
            e = dict(os.environ, **env)
            subprocess.run([sys.executable, __file__, str(N), str(M), str(D)], env=e)
