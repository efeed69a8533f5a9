#!/bin/bash
# per-launch durations of the bench command (cold-cache, serialised): the kernel SHARES of one EM iteration
TAG=$1; WL=${2:-C2}
mkdir -p gpurun_out
B="python bench.py --workload $WL --no-e2e --no-cpu-baseline --no-large --slice-cells 0 --steps 3 --warmup 3"
timeout 400 $B > gpurun_out/${TAG}_${WL}_plain.json 2> gpurun_out/${TAG}_${WL}_plain.err || { echo "plain run failed"; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv -k "regex:^(k_|Device)" --log-file gpurun_out/${TAG}_launches_${WL}.csv $B > gpurun_out/${TAG}_${WL}_ncu.log 2>&1; echo "ncu rc=$?"
python - <<PY
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/${TAG}_launches_${WL}.csv")) if len(r) > 5]
h = rows[0]; ki = h.index("Kernel Name"); vi = h.index("Metric Value")
agg = collections.OrderedDict()
for r in rows[1:]:
    if r[h.index("Metric Name")] != "gpu__time_duration.sum": continue
    a = agg.setdefault(r[ki][:70], [0, 0.0]); a[0] += 1; a[1] += float(r[vi].replace(",", ""))
unit = rows[1][h.index("Metric Unit")]
for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1])[:25]:
    print(f"{n:5d} x {t/n:10.2f} {unit}  {k}")
PY
