#!/bin/bash
# round-end evidence on ONE GPU: tests, the driver's bench line, the reference arm, C5 on one GPU, launch lists, ncu captures
TAG=${1:-rXX}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/${TAG}_gputests.log
CLX_TILED_STATS=1 timeout 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err; echo "reference rc=$?"
timeout 600 python bench.py --workload C5 --no-e2e --no-cpu-baseline > gpurun_out/${TAG}_bench_c5.json 2> gpurun_out/${TAG}_bench_c5.err; echo "C5 rc=$?"
python - <<PY
import json
d = json.load(open("gpurun_out/${TAG}_bench.json"))
print("C2", round(d["ms_per_step"]*1e3,1), "us", "%.3g" % d["value"], {k: round(v,1) for k,v in d["phase_us_per_step"].items()}, "frac", round(d["roofline"]["frac"],3), "e2e", d["e2e"]["seconds_per_job"], "%.3g" % d["e2e"]["value"], "cpu", d["cpu_baseline"]["value"], "parity", d["parity"]["ok"])
l = d["large"]
print("LARGE", l["name"], round(l["ms_per_step"]*1e3,1), "us", "%.3g" % l["value"], {k: round(v,1) for k,v in l["phase_us_per_step"].items()}, "frac", round(l["roofline"]["frac"],3), l["properties"]["ok"], l["parity_slice"]["ok"])
r = json.loads(open("gpurun_out/${TAG}_bench_reference.json").read().strip().splitlines()[-1]); print("REF", r.get("value"), r.get("cpu_baseline"))
c = json.load(open("gpurun_out/${TAG}_bench_c5.json")); print("C5", round(c["ms_per_step"]*1e3,1), "us", "%.3g" % c["value"], {k: round(v,1) for k,v in c["phase_us_per_step"].items()}, c["properties"])
PY
bash tools/launch_list.sh ${TAG} C2 | head -16
bash tools/launch_list.sh ${TAG} C3 | head -16
bash tools/ncu_score.sh ${TAG} C2 C3
