#!/bin/bash
# strong-scaling bench lines of several named workloads on N GPUs: run_multi_cfgs.sh N TAG "C3" "C5" "C5 --equal-cells" ...
N=$1; TAG=$2; shift 2
mkdir -p gpurun_out
i=0
for spec in "$@"; do
  set -- $spec; WL=$1; shift; i=$((i+1))
  out=gpurun_out/${TAG}_n${N}_${WL}_$i
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29520+i)) bench.py --gpus $N --workload $WL "$@" > $out.json 2> $out.err; echo "$spec rc=$?"
  grep -v "^\*\|OMP_NUM\|^$" $out.err | tail -3 | cut -c1-300
  python - <<PY
import json
try:
    d = json.loads(open("$out.json").read().strip().splitlines()[-1])
    print("N=$N $spec", round(d["ms_per_step"]*1e3,1), "us", "%.3g" % d["value"], {k: round(v,1) for k,v in d["phase_us_per_step"].items()}, "frac", round(d["roofline"]["frac"],3))
    for k in ("sharded_parity","speedup_vs_one_gpu_same_matrix","properties"):
        print("  ", k, json.dumps(d.get(k))[:300])
    im = d.get("imbalance") or {}
    print("   imbalance", im.get("nnz_max_over_mean"), im.get("score_us_max_over_mean"), "gen", round(d.get("generate_s",0),2), "load", round(d.get("load_s",0),2))
except Exception as e:
    print("no json:", e)
PY
done
