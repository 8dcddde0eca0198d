#!/bin/bash
# tools/run_one.sh <lib.so> <config> <trajectories>: one bench line of a variant library
lib=$1; cfg=$2; n=$3
MBB200_LIB=$PWD/$lib timeout 600 python bench.py --config $cfg --trajectories $n --steps 1 --warmup 1 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$lib','$cfg', '%.0f traj/s'%d['value'], '%.1f ms'%d['ms_per_step'], 'frac %.4f'%d['roofline']['frac'])"
