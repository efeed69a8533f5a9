#!/bin/bash
# One GPU call: the parity suite, the bench lines (C2 in full, C3 shard), then the ncu evidence for profiles/
# (launch list + ncu --set full of one EM iteration), each ncu pass only after the same command ran clean without it.
TAG=${1:-rXX}
mkdir -p gpurun_out
B="python bench.py"
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/${TAG}_tests.log
timeout 300 $B > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
timeout 300 $B --workload C3s --no-e2e --no-cpu-baseline > gpurun_out/${TAG}_bench_c3shard.json 2> gpurun_out/${TAG}_bench_c3s.err; echo "bench c3s rc=$?"
python - <<PY
import json
for f in ["bench", "bench_c3shard"]:
    try:
        d = json.load(open("gpurun_out/${TAG}_%s.json" % f))
        print(f, round(d["ms_per_step"] * 1e3, 2), "us", "%.3g" % d["value"], d["phase_us_per_step"], round(d["roofline"]["frac"], 4),
              d["e2e"] and d["e2e"]["seconds_per_job"])
    except Exception as e:
        print(f, "unreadable", e)
PY
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_launches.csv \
  $B --no-e2e --no-cpu-baseline --steps 2 --warmup 1 > gpurun_out/${TAG}_ncu1.log 2>&1; echo "launch list rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on \
  -k 'regex:k_score_tiled|k_score_finish|k_tables_em|k_heavy_em|k_select_coop|k_columns|k_locus_filter' -s 15 -c 8 \
  -f -o gpurun_out/${TAG}_iteration $B --no-e2e --no-cpu-baseline --steps 2 --warmup 3 > gpurun_out/${TAG}_ncu2.log 2>&1; echo "full capture rc=$?"
