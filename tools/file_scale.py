"""Developer tool: copy one tools/scale_round.sh visit from gpurun_out/ into profiles/ (bench lines only).
    python tools/file_scale.py <tag> <N>"""
import os, sys
tag, n = sys.argv[1], sys.argv[2]
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for kind in ("weak_c3", "strong_c3", "strong_c5"):
    src = os.path.join(root, "gpurun_out", "%s_%s_n%s.json" % (tag, kind, n))
    if os.path.exists(src):
        lines = [l for l in open(src).read().splitlines() if l.startswith("{")]
        if lines:
            open(os.path.join(root, "profiles", "scale_r02_%s_n%s.json" % (kind, n)), "w").write(lines[-1] + "\n")
src = os.path.join(root, "gpurun_out", "%s_multi_gpu_check_n%s.txt" % (tag, n))
if os.path.exists(src):
    keep = [l for l in open(src).read().splitlines() if l.startswith("world=") or l.startswith("MULTI_GPU")]
    open(os.path.join(root, "profiles", "multi_gpu_check_r02_n%s.txt" % n), "w").write("\n".join(keep) + "\n")
