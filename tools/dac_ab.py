"""Developer tool: k_dac (recomputing) against k_dac3 (kernel values on chip, DVF_DAC_V3=1): results and time."""
import os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1:
    import torch
    from dynamo_release_b200 import kernels
    torch.manual_seed(0)
    outs = {}
    for n, D, M in ((1000000, 30, 500), (100000, 30, 100), (300001, 17, 333)):
        x = torch.randn((n, D), dtype=torch.float64, device="cuda") + 3.0
        Xc = x[:M].clone() + 0.01
        C = torch.randn((M, D), dtype=torch.float64, device="cuda")
        f = lambda: kernels.div_acc_curv(x, Xc, 0.02, C)
        r = f(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            f()
        e1.record(); torch.cuda.synchronize()
        print("%s n=%d D=%d M=%d: %.2f ms" % (sys.argv[1], n, D, M, e0.elapsed_time(e1) / 3), flush=True)
        torch.save({k: v.cpu() for k, v in r.items() if v is not None}, "/tmp/dac_%s_%d.pt" % (sys.argv[1], n))
else:
    import torch
    for tag, env in (("v1", {}), ("v3", {"DVF_DAC_V3": "1"})):
        subprocess.run([sys.executable, __file__, tag], env=dict(os.environ, **env))
    for n in (1000000, 100000, 300001):
        a, b = torch.load("/tmp/dac_v1_%d.pt" % n), torch.load("/tmp/dac_v3_%d.pt" % n)
        print(n, {k: "%.1e" % float((a[k] - b[k]).abs().max() / a[k].abs().max()) for k in a})
