"""Developer tool: time the two control-point solvers in isolation -- dvf_solve (Cholesky, cooperative grid; DVF_SOLVE_GRID)
and dvf_solve_normal (normal equations + LU in one thread-block cluster; DVF_LU_CLUSTER = cluster size) -- and check the
LU result against numpy.linalg.solve of the same normal equations."""
import os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def one(M, D, which):
    import numpy as np, torch
    from dynamo_release_b200 import kernels as k
    rng = np.random.default_rng(0)
    W = k.row_width(M, D)
    A = rng.standard_normal((M + 200, M)) / np.sqrt(M)
    G = np.zeros((W, W)); G[:M, :M] = np.tril(A.T @ A); G[M:M + D, :M] = rng.standard_normal((D, M))
    Kc = np.eye(M)
    Gd, Kd = torch.from_numpy(G).cuda(), torch.from_numpy(Kc).cuda()
    fn = k.solve if which == "chol" else k.solve_normal
    for _ in range(3):
        Cd, info = fn(Gd, Kd, M, D, 0.3)
    lhs = np.tril(G[:M, :M]) + np.tril(G[:M, :M], -1).T + 0.3 * Kc
    rhs = G[M:M + D, :M].T
    ref = np.linalg.solve(lhs.T @ lhs, lhs.T @ rhs)
    err = np.abs(Cd.cpu().numpy() - ref).max() / np.abs(ref).max()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        fn(Gd, Kd, M, D, 0.3)
    e1.record(); torch.cuda.synchronize()
    if which == "lu":   # phase timers of the factor kernel's CTA 0 (kern_solve.cu, LU_MARK)
        import ctypes as C
        from dynamo_release_b200 import _lib
        from dynamo_release_b200.device import lib, ptr, stream_ptr, empty
        nb = C.c_size_t(0)
        lib().dvf_solve_normal_workspace_bytes(M, D, C.byref(nb))
        ws = torch.zeros((nb.value + 7) // 8, dtype=torch.float64, device="cuda")
        Cm = empty((M, D)); inf = torch.zeros(1, dtype=torch.int32, device="cuda")
        _lib.check(lib().dvf_solve_normal(ptr(Gd), Gd.shape[1], ptr(Kd), M, D, 0.3, ptr(ws), nb.value, ptr(Cm), ptr(inf), stream_ptr()))
        torch.cuda.synchronize()
        ldn = (M + D + 1) & ~1
        t = ws.view(torch.int64)[2 * M * ldn + 16: 2 * M * ldn + 24].cpu().numpy() / 1e3
        fine = ws.view(torch.int64)[2 * M * ldn + 24: 2 * M * ldn + 32].cpu().numpy()
        if fine.any():   # built with -DDVF_LU_FINE: cycles of thread 0 per section of the panel's column loop
            print("   panel column loop (cycles per column): elimination %.0f | warp arg-max %.0f | barrier %.0f | scan + pivot row %.0f | barrier %.0f"
                  % tuple(fine[:5] / M), flush=True)
        print("   factor phases (us): panel %.0f | sync %.0f | swaps+U12 %.0f | sync %.0f | trailing %.0f | sync %.0f | back-subst %.0f"
              % tuple(t[:7]), flush=True)
    print("%s M=%d D=%d grid=%s cluster=%s: %.3f ms per solve (incl. ~20 us host wrapper), info %d, rel. dev. from "
          "numpy's normal-equation solve %.2g" % (which, M, D, os.environ.get("DVF_SOLVE_GRID", "auto"),
                                                  os.environ.get("DVF_LU_CLUSTER", "auto"), e0.elapsed_time(e1) / 20, info, err), flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 3:
        one(int(sys.argv[1]), int(sys.argv[2]), sys.argv[3])
    else:
        for M, D in ((500, 30), (1000, 50), (100, 30), (1500, 30)):
            subprocess.run([sys.executable, __file__, str(M), str(D), "chol"])
            for c in ("", "8", "4", "1"):
                env = dict(os.environ)
                if c:
                    env["DVF_LU_CLUSTER"] = c
                subprocess.run([sys.executable, __file__, str(M), str(D), "lu"], env=env)
