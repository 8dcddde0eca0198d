#!/bin/bash
# bench.py --config C5S --links $LINKS once per variant library; prints "name value ms"
N=${N:-8192}; LINKS=${LINKS:-20}; CFG=${CFG:-C5S}
for lib in multibody.jl_b200/lib/variants/lib_*.so; do
  name=$(basename $lib .so)
  MBB200_LIB=$PWD/$lib timeout 600 python bench.py --config $CFG --links $LINKS --trajectories $N --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/cv_$name.json 2> gpurun_out/cv_$name.err
  python - "$name" gpurun_out/cv_$name.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); print(sys.argv[1], "%.1f traj/s"%d["value"], "%.1f ms"%d["ms_per_step"], "frac %.4f"%d["roofline"]["frac"])
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
done
