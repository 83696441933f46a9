#!/bin/bash
# One multi-GPU visit (run under `gpurun --gpus N`): sharded-vs-single parity check, then the bench at N GPUs:
# weak scaling of c3 (what the driver's SCALE run measures), strong scaling of c3 (1M cells in total) and the c5 sweep
# (4M x 50, M = 1000, 10 EM iterations, strong).  Outputs under gpurun_out/ with the given tag.
N=${1:-2}; TAG=${2:-r02}
P=$((29500 + N))
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P "$@"; }
if [ "$N" -gt 1 ]; then
  run tests/multi_gpu_check.py > gpurun_out/${TAG}_multi_gpu_check_n$N.txt 2>&1; echo "check rc=$?"; grep -h "world=\|MULTI_GPU" gpurun_out/${TAG}_multi_gpu_check_n$N.txt
  B="run bench.py"
else
  B="python bench.py"
fi
$B --gpus $N --steps 3 --warmup 3 > gpurun_out/${TAG}_weak_c3_n$N.json 2> gpurun_out/${TAG}_weak_c3_n$N.err; echo "weak rc=$?"
$B --gpus $N --steps 2 --warmup 2 --scaling strong --no-cpu-baseline > gpurun_out/${TAG}_strong_c3_n$N.json 2> gpurun_out/${TAG}_strong_c3_n$N.err; echo "strong rc=$?"
$B --gpus $N --steps 2 --warmup 2 --workload c5 --no-cpu-baseline --no-e2e > gpurun_out/${TAG}_strong_c5_n$N.json 2> gpurun_out/${TAG}_strong_c5_n$N.err; echo "c5 rc=$?"
for f in weak_c3 strong_c3 strong_c5; do python - <<PY
import json
try:
    d = json.loads([l for l in open("gpurun_out/${TAG}_${f}_n$N.json").read().splitlines() if l.startswith("{")][-1])
    print("${f} n=$N value %.4g ms %.1f stages %s e2e %s iters %s" % (d["value"], d["ms_per_step"], {k: round(v, 1) for k, v in d["stages_ms"].items()}, d["e2e"] and round(d["e2e"]["value"]), d["config"]["em_iterations"]))
    if "parity_vs_1gpu" in d: print("   parity_vs_1gpu:", json.dumps(d["parity_vs_1gpu"]))
except Exception as e:
    print("${f}: no line", e)
PY
done
