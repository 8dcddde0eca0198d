#!/bin/bash
# Round-2 measurement batch (run under gpurun on one B200): GPU suite, bench lines of every config, launch list, and one
# `ncu --set full` capture per benchmarked kernel (metrics extracted here; only the robot report travels back).
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -n 15 > $O/r02_gputest_final.log
b() { name=$1; shift; timeout 900 python bench.py "$@" > $O/r02_bench_$name.json 2> $O/r02_bench_$name.err; tail -c 400 $O/r02_bench_$name.err | tail -n 2; }
b C4 --steps 1 --warmup 1
b C2B --config C2B --steps 1 --warmup 1 --cpu-sample 2048
b C3 --config C3 --steps 1 --warmup 1 --cpu-sample 2048
b C3P --config C3P --steps 1 --warmup 1 --cpu-sample 2048
b C5 --config C5 --steps 1 --warmup 1 --cpu-sample 64
b C5E --config C5E --steps 1 --warmup 1 --cpu-sample 64
b C5S_links20 --config C5S --links 20 --steps 1 --warmup 1 --no-e2e
b C5_links20 --config C5 --links 20 --trajectories 8192 --steps 1 --warmup 1 --no-e2e
b C5S_links50 --config C5S --links 50 --trajectories 8192 --steps 1 --warmup 1 --no-e2e
b C5_links50 --config C5 --links 50 --trajectories 1184 --steps 1 --warmup 0 --no-e2e --no-cpu-baseline
# launch list of the headline command
ncu --metrics gpu__time_duration.sum --clock-control none -c 20 --csv --log-file $O/r02_launches_v9.csv \
    python bench.py --trajectories 37888 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > $O/r02_launches_v9.log 2>&1
RX='dram__bytes_(read|write)\.sum|gpu__time_duration\.sum|pipe_fp64|registers_per_thread|local_(ld|st)|sector_hit_rate\.pct|warps_active\.avg|issue_stalled_(long_scoreboard|barrier|no_instruction|wait|short_scoreboard|math_pipe_throttle|lg_throttle)_per_warp_active|smsp__inst_executed\.sum |launch__(block_size|grid_size|shared_mem_per_block_dynamic)|sm__throughput|gpu__dram_throughput'
cap() { name=$1; pat=$2; shift 2
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$pat -c 1 -f -o $O/prof_r02_$name python bench.py "$@" --steps 1 --warmup 0 --no-cpu-baseline --no-e2e > $O/prof_r02_$name.log 2>&1
  ncu -i $O/prof_r02_$name.ncu-rep --page raw --csv > $O/prof_r02_${name}_raw_full.csv 2>/dev/null
  python - "$O/prof_r02_${name}_raw_full.csv" "$O/prof_r02_${name}_raw.txt" <<'PY'
import csv, re, sys
rx = re.compile(r'dram__bytes_(read|write)\.sum$|gpu__time_duration\.sum|pipe_fp64|registers_per_thread|local_(ld|st)|sector_hit_rate\.pct|warps_active\.avg|issue_stalled_.*_per_warp_active|^smsp__inst_executed\.sum$|launch__(block_size|grid_size|shared_mem_per_block_dynamic|occupancy_limit)|sm__throughput\.avg\.pct|gpu__dram_throughput\.avg\.pct|^Kernel Name|lts__t_bytes\.sum$|l1tex__data_pipe|inst_executed_pipe_fp64')
try:
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units, vals = rows[0], rows[1], rows[2]
    with open(sys.argv[2], "w") as f:
        for h, u, v in zip(hdr, units, vals):
            if rx.search(h):
                f.write(f"{h} [{u}] = {v}\n")
except Exception as e:
    open(sys.argv[2], "w").write(f"extraction failed: {e}\n")
PY
  rm -f $O/prof_r02_${name}_raw_full.csv
  [ "$name" != "robot_v9" ] && rm -f $O/prof_r02_$name.ncu-rep
}
cap robot_v9 rosenbrock_rows --trajectories 37888
cap C2 ensemble_kernel --config C2 --trajectories 151552
cap C2B ensemble_kernel --config C2B --trajectories 37888
cap C3 ensemble_kernel --config C3 --trajectories 37888
cap C5 rosenbrock_rows --config C5 --trajectories 4736
cap C5E rosenbrock_rows --config C5E --trajectories 4736
cap C5S_links20 chain_warp --config C5S --links 20 --trajectories 1184
ls -la $O | tail -n 40
cat $O/r02_gputest_final.log
