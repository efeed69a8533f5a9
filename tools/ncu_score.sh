#!/bin/bash
# one ncu --set full capture of k_score_runs per workload given (after the plain command exited 0)
TAG=$1; shift
mkdir -p gpurun_out
for WL in "$@"; do
  B="python bench.py --workload $WL --no-e2e --no-cpu-baseline --slice-cells 0 --steps 5"
  timeout 400 $B > gpurun_out/${TAG}_${WL}_plain.json 2> gpurun_out/${TAG}_${WL}_plain.err || { echo "plain run failed for $WL"; continue; }
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_score_runs -s 4 -c 1 -f -o gpurun_out/${TAG}_score_${WL} $B > gpurun_out/${TAG}_${WL}_ncu.log 2>&1; echo "$WL ncu rc=$?"
done
