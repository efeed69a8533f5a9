#!/bin/bash
# tests, C3s bench, then one ncu --set full capture of the scoring kernel on C3s
TAG=${1:-rXX}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/${TAG}_tests.log
export CLX_TILED_STATS=1
B="python bench.py --workload C3s --no-e2e --no-cpu-baseline --steps 5"
timeout 400 $B > gpurun_out/${TAG}_bench_c3s.json 2> gpurun_out/${TAG}_bench_c3s.err && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_score_runs -s 4 -c 1 -f -o gpurun_out/${TAG}_score_c3s $B > gpurun_out/${TAG}_ncu.log 2>&1; echo "ncu rc=$?"
grep "clx" gpurun_out/${TAG}_bench_c3s.err | head -3
python - <<PY
import json
for f in ["bench_c3s"]:
    try:
        d = json.load(open("gpurun_out/${TAG}_%s.json" % f))
        print(f, round(d["ms_per_step"] * 1e3, 2), "us", "%.3g" % d["value"], {k: round(v, 1) for k, v in d["phase_us_per_step"].items()},
              "frac", round(d["roofline"]["frac"], 4))
    except Exception as e:
        print(f, "unreadable", e)
PY
