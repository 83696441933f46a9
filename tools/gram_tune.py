"""Developer tool: time the Gram kernel (dvf_gram) alone for different schedule knobs (env DVF_GRAM_*)."""
import os, sys, subprocess, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

def one(N, M, D):
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
    e0.record()
    for _ in range(5):
        G = kernels.gram(UY, P, 1e-5, D, M)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    flops = N * (M * M + M + 2.0 * M * D)
    print("N=%d M=%d D=%d  %.3f ms  %.2f TF algorithmic  env=%s" % (N, M, D, ms, flops / ms / 1e9,
          {k: v for k, v in os.environ.items() if k.startswith("DVF_GRAM")}), flush=True)

if __name__ == "__main__":
    if len(sys.argv) > 1:
        one(int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]))
    else:
        runs = [({}, (1000000, 500, 30))]
        for G in (4, 7, 10, 11, 12, 14, 18, 28):
            runs.append(({"DVF_GRAM_G": str(G)}, (1000000, 500, 30)))
        for sb in (8, 16, 64, 128):
            runs.append(({"DVF_GRAM_SB": str(sb)}, (1000000, 500, 30)))
        runs.append(({}, (1000000, 482, 30)))     # 512 columns: FULL + HALF only
        runs.append(({}, (1000000, 100, 30)))
        runs.append(({}, (1000000, 1000, 50)))
        for env, (N, M, D) in runs:
            e = dict(os.environ, DVF_GRAM_DEBUG="1", **env)
            subprocess.run([sys.executable, __file__, str(N), str(M), str(D)], env=e)
