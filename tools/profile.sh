#!/bin/bash
# One GPU call: the ncu evidence for profiles/ (run only after the same commands exited 0 without ncu).
#   1. launch list of a short bench run (per-launch gpu__time_duration, cold-cache and serialised)
#   2. ncu --set full of one EM iteration's kernels (third iteration of the run)
# plus a few one-line A/B bench runs of tunables (no ncu).
TAG=${1:-r01p}
mkdir -p gpurun_out
B="python bench.py --no-e2e --no-cpu-baseline"
timeout 200 $B --steps 2 --warmup 1 > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.err; exit 1; }
for v in 1 3 4; do
  CLX_TILED_ITEMS_PER_SM=$v timeout 200 $B > gpurun_out/${TAG}_items$v.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_items$v.json")); print("items/SM $v:", d["ms_per_step"], d["phase_us_per_step"])
PY
done
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_launches.csv \
  $B --steps 2 --warmup 1 > gpurun_out/${TAG}_ncu1.log 2>&1
echo "launch list rc=$?"
timeout 500 ncu --set full --clock-control none --import-source on \
  -k 'regex:k_score_tiled|k_score_finish|k_tables_em|k_heavy_em|k_select_coop|k_columns|k_locus_filter' -s 15 -c 8 \
  -f -o gpurun_out/${TAG}_iteration $B --steps 2 --warmup 3 > gpurun_out/${TAG}_ncu2.log 2>&1
echo "full capture rc=$?"; ls -la gpurun_out/ | tail -12
