#!/bin/bash
# L2 set-aside experiment on the robot kernel: base vs LU slots in global scratch, without / with the access-policy window
N=${N:-151552}
V=multibody.jl_b200/lib/variants
run() { name=$1; shift; env "$@" timeout 300 python bench.py --trajectories $N --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/l2_$name.json 2> gpurun_out/l2_$name.err
  python - "$name" gpurun_out/l2_$name.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); print(sys.argv[1], "%.0f traj/s"%d["value"], "%.1f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"])
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
run glu MBB200_LIB=$PWD/$V/lib_glu.so
run glu_persist MBB200_LIB=$PWD/$V/lib_glu.so MBB200_L2_PERSIST=1 MBB200_L2_DEBUG=1
run glu_persist24 MBB200_LIB=$PWD/$V/lib_glu.so MBB200_L2_PERSIST=24 MBB200_L2_DEBUG=1
grep mbb200: gpurun_out/l2_glu_persist.err | head -2
MBB200_LIB=$PWD/$V/lib_glu.so MBB200_L2_PERSIST=1 timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "config_parity and C4" 2>&1 | tail -2
