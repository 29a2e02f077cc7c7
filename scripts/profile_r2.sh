#!/bin/bash
# Round-2 profiles (run on the GPU box through gpurun; every ncu pass only after the same command exited 0 without ncu).
# 1. launch list of the learn benchmark (gpu__time_duration per launch)   2. ncu --set full of the persistent loop kernel
# 3. ncu --set full of the pair-count kernel on the 46 M-token probe
set -u
OUT=gpurun_out
B="python bench.py --steps 400 --warmup 5 --no-induce --no-medoid --no-cpu --no-e2e"
$B > $OUT/r2_prof_bench.json 2> $OUT/r2_prof_bench.err || { echo "bench failed"; tail -5 $OUT/r2_prof_bench.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/r2_launches.csv $B > $OUT/r2_ncu_launches.log 2>&1
echo "launch list rc $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:h2_run -c 1 -o $OUT/r2_h2run_full -f $B --no-scale-probe > $OUT/r2_ncu_h2.log 2>&1
echo "h2_run full rc $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_slot_count -s 14 -c 1 -o $OUT/r2_k1_scale_full -f $B > $OUT/r2_ncu_k1.log 2>&1
echo "k_slot_count full rc $?"
for f in r2_h2run_full r2_k1_scale_full; do
  [ -f $OUT/$f.ncu-rep ] && ncu -i $OUT/$f.ncu-rep --page raw --csv > $OUT/${f}_raw.csv 2>/dev/null
done
ls -la $OUT | grep r2_
