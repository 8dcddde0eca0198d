#!/bin/bash
# robot kernel v12 (rows mode 6, 192 threads): GPU suite, full bench line, launch list, one ncu --set full capture
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -n 6 > $O/r02_gputest_v12.log
timeout 900 python bench.py > $O/r02_bench_C4_v12.json 2> $O/r02_bench_C4_v12.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 20 --csv --log-file $O/r02_launches_v12.csv \
    python bench.py --trajectories 28416 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > $O/r02_launches_v12.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:rosenbrock_rows -c 1 -f -o $O/prof_r02_robot_v12 python bench.py --trajectories 28416 --steps 1 --warmup 0 --no-cpu-baseline --no-e2e > $O/prof_r02_robot_v12.log 2>&1
ncu -i $O/prof_r02_robot_v12.ncu-rep --page raw --csv > $O/prof_r02_robot_v12_raw_full.csv 2>/dev/null
python - "$O/prof_r02_robot_v12_raw_full.csv" "$O/r02_ncu_robot_v12.txt" <<'PY'
import csv, re, sys
rx = re.compile(r'dram__bytes_(read|write)\.sum$|gpu__time_duration\.sum|pipe_fp64|registers_per_thread|local_(ld|st)|sector_hit_rate\.pct|warps_active\.avg|issue_stalled_.*_per_issue_active|^smsp__inst_executed\.sum$|launch__(block_size|grid_size|shared_mem_per_block_dynamic|occupancy_limit)|sm__throughput\.avg\.pct|gpu__dram_throughput\.avg\.pct|^Kernel Name|lts__t_bytes\.sum$|l1tex__data_pipe_lsu_wavefronts(_mem_shared)?\.sum|inst_executed_pipe_fp64|smsp__issue_active.avg.pct')
rows = list(csv.reader(open(sys.argv[1])))
hdr, units, vals = rows[0], rows[1], rows[2]
with open(sys.argv[2], "w") as f:
    for h, u, v in zip(hdr, units, vals):
        if rx.search(h):
            f.write(f"{h} [{u}] = {v}\n")
PY
rm -f $O/prof_r02_robot_v12_raw_full.csv
cat $O/r02_gputest_v12.log; tail -c 300 $O/r02_bench_C4_v12.json
