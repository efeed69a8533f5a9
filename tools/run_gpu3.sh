#!/bin/bash
# ncu capture of the scorer on C3s, then the default bench line (C2 headline + large C3)
TAG=${1:-rXX}
mkdir -p gpurun_out
export CLX_TILED_STATS=1
B="python bench.py --workload C3s --no-e2e --no-cpu-baseline --steps 5"
timeout 400 $B > gpurun_out/${TAG}_bench_c3s.json 2> gpurun_out/${TAG}_bench_c3s.err && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_score_runs -s 4 -c 1 -f -o gpurun_out/${TAG}_score_c3s $B > gpurun_out/${TAG}_ncu.log 2>&1; echo "ncu rc=$?"
timeout 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
tail -5 gpurun_out/${TAG}_bench.err
python - <<PY
import json
d = json.load(open("gpurun_out/${TAG}_bench.json"))
print("C2", round(d["ms_per_step"]*1e3,1), "us", {k: round(v,1) for k,v in d["phase_us_per_step"].items()}, "frac", round(d["roofline"]["frac"],3), "e2e", d["e2e"] and d["e2e"]["seconds_per_job"])
print("parity", d.get("parity"))
l = d.get("large")
if l:
    print("LARGE", l["name"], round(l["ms_per_step"]*1e3,1), "us", "%.3g" % l["value"], {k: round(v,1) for k,v in l["phase_us_per_step"].items()}, "frac", round(l["roofline"]["frac"],3), "gen", round(l["generate_s"],2), "load", round(l["load_s"],2))
    print(l["properties"], l.get("parity_slice"))
PY
