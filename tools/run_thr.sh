#!/bin/bash
# threads per CTA of the robot kernel at a size where the rounds do not matter (606 208 = 16 x 37 888)
for t in 256 224 192; do
  MBB200_THREADS_PER_CTA=$t timeout 300 python bench.py --trajectories 606208 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/thr_$t.json 2> gpurun_out/thr_$t.err
  python - "$t" gpurun_out/thr_$t.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); print(sys.argv[1], "%.0f traj/s"%d["value"], "%.1f ms"%d["ms_per_step"])
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
done
