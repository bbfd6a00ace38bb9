#!/bin/bash
# Launch list of one bench configuration: every kernel of a few steady-state steps with its device time
# (serialised, cold caches: compare SHARES).  usage: prof_launches.sh <name> <bench args...>   -> gpurun_out/prof/<name>.csv
name=$1; shift
out=gpurun_out/prof; mkdir -p $out
python bench.py --steps 4 --warmup 3 --no-e2e --no-baseline --min-timed-ms 1 "$@" > $out/$name.plain.json 2> $out/$name.plain.err || { echo "plain run failed"; tail -5 $out/$name.plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,launch__registers_per_thread,launch__grid_size,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active \
    --clock-control none --csv --log-file $out/$name.csv \
    python bench.py --steps 4 --warmup 3 --no-e2e --no-baseline --min-timed-ms 1 "$@" > $out/$name.ncu.log 2>&1
echo "ncu exit $?"; wc -l $out/$name.csv
