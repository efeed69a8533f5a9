#!/bin/bash
# A/B sweep of scorer layout knobs on one workload: prints the score-phase time per setting
WL=${1:-C3s}; shift
mkdir -p gpurun_out
run() {
  out=$(env "$@" CLX_TILED_STATS=1 timeout 300 python bench.py --workload $WL --no-e2e --no-cpu-baseline --slice-cells 0 --steps 10 2> gpurun_out/sweep.err)
  echo "$@" "$(echo "$out" | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print({k:round(v,1) for k,v in d['phase_us_per_step'].items()},round(d['ms_per_step']*1e3,1),'us/iter score',round(d['phase_us_per_step']['score'],1),'frac',round(d['roofline']['frac'],3))")" "$(grep -o 'NB [0-9]* W [0-9]* Pc [0-9]*' gpurun_out/sweep.err | head -1)" "$(grep -o 'chunks.*' gpurun_out/sweep.err | head -1)"
}
for setting in "$@"; do run $setting; done
