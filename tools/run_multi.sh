#!/bin/bash
# multi-GPU: sharded parity tests, then the strong-scaling bench line on N GPUs
N=${1:-2}; TAG=${2:-rXX}; shift 2
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/${TAG}_multi_tests.log
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" > gpurun_out/${TAG}_n${N}.json 2> gpurun_out/${TAG}_n${N}.err; echo "bench rc=$?"
tail -5 gpurun_out/${TAG}_n${N}.err | cut -c1-300
python - <<PY
import json
d = json.loads(open("gpurun_out/${TAG}_n${N}.json").read().strip().splitlines()[-1])
print(d["config"].get("workload","")[:80])
print("N=$N", round(d["ms_per_step"]*1e3,1), "us", "%.3g" % d["value"], {k: round(v,1) for k,v in d["phase_us_per_step"].items()}, "frac", round(d["roofline"]["frac"],3))
for k in ("sharded_parity","same_matrix_on_one_gpu","speedup_vs_one_gpu_same_matrix","imbalance","e2e"):
    print(k, json.dumps(d.get(k))[:400])
PY
