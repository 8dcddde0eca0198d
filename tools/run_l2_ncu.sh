#!/bin/bash
# one ncu capture of the robot kernel with the LU slots in the L2 set-aside (tools/run_l2.sh is the timing half)
O=gpurun_out; V=$PWD/multibody.jl_b200/lib/variants
for mode in persist plain; do
  if [ $mode = persist ]; then export MBB200_L2_PERSIST=1; else unset MBB200_L2_PERSIST; fi
  MBB200_LIB=$V/lib_glu.so timeout 600 ncu --set full --clock-control none -k regex:rosenbrock_rows -c 1 -f -o $O/prof_glu_$mode python bench.py --trajectories 37888 --steps 1 --warmup 0 --no-cpu-baseline --no-e2e > $O/prof_glu_$mode.log 2>&1
  ncu -i $O/prof_glu_$mode.ncu-rep --page raw --csv > $O/prof_glu_${mode}_raw.csv 2>/dev/null
  rm -f $O/prof_glu_$mode.ncu-rep
done
