"""Developer tool: turn one tools/gpu_round.sh run (gpurun_out/*_<tag>*) into the tracked files under profiles/.

  python tools/profile_collect.py r01g [out-suffix]

writes profiles/launches_<sfx>.csv + launches_<sfx>_summary.txt (per-kernel totals and shares of the ncu launch
list), profiles/ncu_<kernel>_<sfx>_details.txt (ncu --page details of the --set full capture, plus the raw DRAM
byte counters and pipe utilisation), and copies the bench lines."""
import collections, csv, io, os, shutil, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
PROF = os.path.join(ROOT, "profiles")
RAW_KEYS = ["dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum",
            "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
            "sm__pipe_shared_cycles_active.avg.pct_of_peak_sustained_active",
            "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_uniform.sum", "lts__t_sector_hit_rate.pct",
            "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "launch__registers_per_thread",
            "launch__occupancy_limit_registers", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "smsp__cycles_active.avg", "sm__throughput.avg.pct_of_peak_sustained_elapsed"]


def launch_summary(tag, sfx):
    src = os.path.join(OUT, "launches_%s.csv" % tag)
    rows = []
    with open(src) as f:
        lines = [l for l in f if not l.startswith("==")]
    rd = csv.DictReader(io.StringIO("".join(lines)))
    for r in rd:
        if r.get("Metric Name") == "gpu__time_duration.sum":
            v = float(r["Metric Value"].replace(",", ""))
            unit = r.get("Metric Unit", "ns")
            v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(unit, 1e-6)
            rows.append((r["Kernel Name"], v))
    # the FP64 peak microbenchmark bench.py runs before the timed region (tools/fp64_peak) is not part of the step
    micro = [r for r in rows if r[0].split("(")[0].strip().split("<")[0].replace("void ", "") in ("k_dmma16816", "k_dmma884", "k_dfma", "k_exp")]
    rows = [r for r in rows if r not in micro]
    tot = sum(v for _, v in rows)
    agg = collections.OrderedDict()
    for k, v in rows:
        k = k.split("(")[0]
        a = agg.setdefault(k, [0.0, 0])
        a[0] += v
        a[1] += 1
    shutil.copy(src, os.path.join(PROF, "launches_%s.csv" % sfx))
    with open(os.path.join(PROF, "launches_%s_summary.txt" % sfx), "w") as f:
        f.write("ncu --metrics gpu__time_duration.sum --clock-control none : python bench.py --steps 1 --warmup 1 "
                "--no-e2e --no-cpu-baseline  (gpurun tag %s; %d launches, %.3f ms total)\n" % (tag, len(rows), tot))
        f.write("cold-cache, serialised launches: the SHARE is what compares with bench.py's CUDA-event timing\n")
        f.write("(%d launches of the in-run FP64 peak microbenchmark, %.1f ms, left out)\n\n" % (len(micro), sum(v for _, v in micro)))
        for k, (v, c) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
            f.write("%9.3f ms %5.1f%% x%-4d %s\n" % (v, 100 * v / tot, c, k[:90]))
    print("launch list: %d launches, %.1f ms" % (len(rows), tot))


def ncu_details(tag, sfx):
    for fn in sorted(os.listdir(OUT)):
        if not fn.endswith("_%s.ncu-rep" % tag):
            continue
        k = fn[:-len("_%s.ncu-rep" % tag)]
        rep = os.path.join(OUT, fn)
        det = subprocess.run(["ncu", "-i", rep, "--page", "details"], capture_output=True, text=True).stdout
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        picked = []
        try:
            rr = list(csv.reader(io.StringIO(raw)))
            hdr, units, vals = rr[0], rr[1], rr[2]
            for key in RAW_KEYS:
                if key in hdr:
                    i = hdr.index(key)
                    picked.append("%-70s %-12s %s" % (key, units[i], vals[i]))
        except Exception as e:  # a capture that matched no kernel
            picked.append("raw page unreadable: %r" % (e,))
        with open(os.path.join(PROF, "ncu_%s_%s_details.txt" % (k, sfx)), "w") as f:
            f.write("ncu --set full --clock-control none --import-source on -k regex:^%s$ --launch-skip 1 -c 1 : "
                    "python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline  (gpurun tag %s)\n\n" % (k, tag))
            f.write("selected raw counters (per launch)\n" + "\n".join(picked) + "\n\n")
            f.write(det)
        print("ncu details:", k, "|", (picked[0] if picked else ""))


if __name__ == "__main__":
    tag = sys.argv[1]
    sfx = sys.argv[2] if len(sys.argv) > 2 else "r01"
    launch_summary(tag, sfx)
    ncu_details(tag, sfx)
    for a, b in (("bench_%s.json" % tag, "bench_%s_n1.json" % sfx), ("bench_ref_%s.json" % tag, "bench_%s_reference.json" % sfx)):
        p = os.path.join(OUT, a)
        if os.path.exists(p) and os.path.getsize(p):
            shutil.copy(p, os.path.join(PROF, b))
