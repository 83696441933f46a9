"""Developer tool: stall samples per CUDA source line of an ncu --set full --import-source on capture.
    python tools/ncu_lines.py gpurun_out/<kernel>_<tag>.ncu-rep [top]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
agg = {}
ci = None
for r in rows:
    if len(r) > 6 and r[0] == "Line No":
        ci = r.index("# Samples"); ii = r.index("Instructions Executed"); continue
    if ci is None or len(r) <= ci or not r[0].strip().isdigit():
        continue
    try:
        s = int(r[ci]); n = int(r[ii])
    except ValueError:
        continue
    a = agg.setdefault(int(r[0]), [0, 0, r[1]]); a[0] += s; a[1] += n
tot = sum(v[0] for v in agg.values()) or 1
print("total samples", tot)
for ln, (s, n, src) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%5d %6d %5.1f%% inst %10d | %s" % (ln, s, 100 * s / tot, n, src.strip()[:120]))
