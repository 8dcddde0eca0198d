#!/bin/bash
# Runs bench.py once per variant library; prints "name value ms".
N=${N:-75776}
for lib in multibody.jl_b200/lib/variants/lib_*.so; do
  name=$(basename $lib .so)
  MBB200_LIB=$PWD/$lib timeout 300 python bench.py --trajectories $N --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/var_$name.json 2> gpurun_out/var_$name.err
  python - "$name" gpurun_out/var_$name.json <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[2])); print(sys.argv[1], "%.0f traj/s"%d["value"], "%.1f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"])
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
done
