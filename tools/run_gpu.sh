#!/bin/bash
# One GPU call: parity suite, then bench lines (C2 headline, C3 shard), results into gpurun_out/<tag>_*
TAG=${1:-rXX}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/${TAG}_tests.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
timeout 400 python bench.py --workload C3s --no-e2e --no-cpu-baseline > gpurun_out/${TAG}_bench_c3s.json 2> gpurun_out/${TAG}_bench_c3s.err; echo "bench c3s rc=$?"
python - <<PY
import json
for f in ["bench", "bench_c3s"]:
    try:
        d = json.load(open("gpurun_out/${TAG}_%s.json" % f))
        print(f, round(d["ms_per_step"] * 1e3, 2), "us", "%.3g" % d["value"], {k: round(v, 1) for k, v in d["phase_us_per_step"].items()},
              "frac", round(d["roofline"]["frac"], 4), "e2e", d["e2e"] and d["e2e"]["seconds_per_job"])
    except Exception as e:
        print(f, "unreadable", e)
PY
tail -3 gpurun_out/${TAG}_bench.err gpurun_out/${TAG}_bench_c3s.err
