#!/bin/bash
# L2-residency probe: cap the resident CTAs per SM for a given variant library
N=${N:-37888}
for spec in "$@"; do
  name="${spec%%:*}"; cap="${spec#*:}"
  MBB200_CTAS_PER_SM=$cap MBB200_LIB=$PWD/multibody.jl_b200/lib/variants/lib_$name.so timeout 300 python bench.py --trajectories $N --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/cap_${name}_$cap.json 2> gpurun_out/cap_${name}_$cap.err
  python - "$name cap=$cap" gpurun_out/cap_${name}_$cap.json <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[2])); print(sys.argv[1], "%.0f traj/s"%d["value"], "%.1f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"])
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
done
