#!/bin/bash
# One GPU call: parity suite on the current build, then the bench lines; on a parity failure, diagnose with the
# debug build and validate/bench the previous build instead (build_prev/ holds both, see CLX_LIB_PATH in api.py).
mkdir -p gpurun_out
PREV=$PWD/build_prev/libcellector_cuda_prev.so
DBG=$PWD/build_prev/libcellector_cuda_dbg.so
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/new_tests.log 2>&1
rc=$?
echo "new_tests rc=$rc"; tail -3 gpurun_out/new_tests.log
if [ $rc -ne 0 ]; then
  CLX_LIB_PATH=$DBG timeout 120 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "small" > gpurun_out/dbg.log 2>&1
  echo "dbg rc=$?"; grep -m 20 "bad " gpurun_out/dbg.log
  export CLX_LIB_PATH=$PREV
  timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/prev_tests.log 2>&1
  echo "prev_tests rc=$?"; tail -3 gpurun_out/prev_tests.log
fi
timeout 300 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench rc=$?"; cat gpurun_out/bench.json
timeout 300 python bench.py --workload C3s --no-e2e --no-cpu-baseline > gpurun_out/bench_c3s.json 2> gpurun_out/bench_c3s.err
echo "bench c3s rc=$?"; cat gpurun_out/bench_c3s.json
