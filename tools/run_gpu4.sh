#!/bin/bash
# full GPU test suite, then the default bench line (C2 headline + large C3)
TAG=${1:-rXX}
mkdir -p gpurun_out
export CLX_TILED_STATS=1
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/${TAG}_tests.log
timeout 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
grep "run stream" gpurun_out/${TAG}_bench.err | sort -u | cut -c1-250
python - <<PY
import json
d = json.load(open("gpurun_out/${TAG}_bench.json"))
print("C2", round(d["ms_per_step"]*1e3,1), "us", {k: round(v,1) for k,v in d["phase_us_per_step"].items()}, "frac", round(d["roofline"]["frac"],3), "e2e", d["e2e"] and d["e2e"]["seconds_per_job"])
print("parity", d.get("parity", {}).get("ok"))
l = d.get("large")
if l:
    print("LARGE", l["name"], round(l["ms_per_step"]*1e3,1), "us", "%.3g" % l["value"], {k: round(v,1) for k,v in l["phase_us_per_step"].items()}, "frac", round(l["roofline"]["frac"],3), "gen", round(l["generate_s"],2), "load", round(l["load_s"],2))
    print(l["properties"], l.get("parity_slice", {}).get("ok"))
PY
