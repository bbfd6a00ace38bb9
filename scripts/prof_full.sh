#!/bin/bash
# `ncu --set full` rows for every kernel of ONE timed window of each bench configuration (1 GPU, never multi-rank).
# Plain run first (must exit 0), then ONE ncu run over all configs in a single process.  Only the CSV export travels
# back (the .ncu-rep of ~50 full-set captures exceeds gpurun's 64 MiB return limit and is deleted on the box).
# usage: prof_full.sh [configs...]   -> gpurun_out/prof/full_all_raw.csv
cfgs=${@:-orr pa fgm kmeans mcpa}
out=gpurun_out/prof; mkdir -p $out
python scripts/prof_full_run.py $cfgs > $out/full_plain.log 2>&1 || { echo "plain run failed"; tail -20 $out/full_plain.log; exit 1; }
OMLDM_PROFILE_WINDOW=1 ncu --set full --clock-control none --profile-from-start off \
    -f -o /tmp/full_all python scripts/prof_full_run.py $cfgs > $out/full_ncu.log 2>&1
echo "ncu exit $?"
ncu -i /tmp/full_all.ncu-rep --page raw --csv > $out/full_all_raw.csv 2>/dev/null
ls -la /tmp/full_all.ncu-rep; wc -l $out/full_all_raw.csv
