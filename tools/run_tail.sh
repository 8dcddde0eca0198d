#!/bin/bash
# strong-scaling tail: per-GPU shards of the 8- and 4-GPU runs with the balanced thread count (default) and with full CTAs
for N in 131072 262144; do for t in auto 256; do
  if [ $t = auto ]; then unset MBB200_THREADS_PER_CTA; else export MBB200_THREADS_PER_CTA=$t; fi
  timeout 300 python bench.py --trajectories $N --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/tail_${N}_$t.json 2> gpurun_out/tail_${N}_$t.err
  python - "$N x $t" gpurun_out/tail_${N}_$t.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); print(sys.argv[1], "%.0f traj/s"%d["value"], "%.1f ms"%d["ms_per_step"])
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
done; done
