"""Developer tool: time the control-point solve (dvf_solve) for different cooperative grid sizes (DVF_SOLVE_GRID)."""
import os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def one(M, D):
    import numpy as np, torch
    from dynamo_release_b200 import kernels as k
    rng = np.random.default_rng(0)
    W = k.row_width(M, D)
    A = rng.standard_normal((M + 200, M)) / np.sqrt(M)
    G = np.zeros((W, W)); G[:M, :M] = np.tril(A.T @ A); G[M:M + D, :M] = rng.standard_normal((D, M))
    Kc = np.eye(M)
    Gd, Kd = torch.from_numpy(G).cuda(), torch.from_numpy(Kc).cuda()
    for _ in range(3):
        k.solve(Gd, Kd, M, D, 0.3)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        k.solve(Gd, Kd, M, D, 0.3)
    e1.record(); torch.cuda.synchronize()
    print("M=%d D=%d grid=%s: %.3f ms per solve (incl. ~20 us host wrapper)" % (M, D, os.environ.get("DVF_SOLVE_GRID", "auto"), e0.elapsed_time(e1) / 20), flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 2:
        one(int(sys.argv[1]), int(sys.argv[2]))
    else:
        for M, D in ((500, 30), (1000, 50), (100, 30)):
            for g in ("", "16", "24", "36", "40", "48", "64", "96", "148"):
                env = dict(os.environ)
                if g:
                    env["DVF_SOLVE_GRID"] = g
                subprocess.run([sys.executable, __file__, str(M), str(D)], env=env)
