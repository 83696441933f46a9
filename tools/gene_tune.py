"""Developer tool: time the gene-space transform Q_e J Q_r^T (C4 shape: D = 50, 2000 x 2000 genes) on a cell slab, for the
tensor-core path and (DVF_GENE_GEMM=0, subprocess) the per-effector kernel; full float32 output and group-sum-only mode."""
import os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def one(n, D, Ge, Gr, mode):
    import torch
    from dynamo_release_b200 import kernels
    torch.manual_seed(0)
    J = torch.randn((D, D, n), dtype=torch.float64, device="cuda")
    Qe = torch.randn((Ge, D), dtype=torch.float64, device="cuda") / 7
    Qr = torch.randn((Gr, D), dtype=torch.float64, device="cuda") / 7
    lab = (torch.arange(n, device="cuda") % 10).to(torch.int32)
    kw = dict(full=True) if mode == "full" else dict(full=False, take_abs=True, group=lab, n_groups=10)

    def run():
        return kernels.gene_transform(J, Qe, Qr, **kw)
    out = run(); del out
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = run()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    flops = 2.0 * Ge * D * (D + Gr) * n
    err = float("nan")
    if mode == "full":
        idx = torch.tensor([0, 1, n // 2, n - 1], device="cuda")
        ref = torch.einsum("ei,ijc,rj->erc", Qe, J[:, :, idx], Qr)
        err = float((out[0][:, :, idx].double() - ref).abs().max() / ref.abs().max())
    print("gene %s n=%d D=%d Ge=%d Gr=%d  %.2f ms  %.2f TF  %.0f cells/s  relerr(f32)=%.1e  gemm=%s" % (
        mode, n, D, Ge, Gr, ms, flops / ms / 1e9, n / ms * 1e3, err, os.environ.get("DVF_GENE_GEMM", "auto")), flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 5:
        one(int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), sys.argv[5])
    else:
        for cfg in [(2048, 50, 2000, 2000, "full"), (1024, 50, 2000, 2000, "group"), (8192, 30, 500, 500, "full")]:
            for gemm in ("1", "0"):
                subprocess.run([sys.executable, __file__] + [str(c) for c in cfg], env=dict(os.environ, DVF_GENE_GEMM=gemm))
