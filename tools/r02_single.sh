#!/bin/bash
# Round-2 single-GPU session: tests, bench, launch list, ncu --set full of every kernel.
# SKIP_TESTS=1 skips pytest. Only CSV exports travel back (the .ncu-rep is too large for the
# 64 MiB return channel and is deleted on the box after export).
OUT=gpurun_out; mkdir -p $OUT
if [ -z "$SKIP_TESTS" ]; then
  timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > $OUT/r02_gputest.log; cat $OUT/r02_gputest.log
fi
timeout 300 python bench.py --steps 20 --warmup 5 --no-configs > $OUT/r02_bench_noconf.json 2> $OUT/r02_bench_noconf.err; echo "bench rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/r02_launches.csv python bench.py --steps 20 --warmup 5 --no-configs --no-cpu > $OUT/r02_ncu_bench.log 2>&1; echo "ncu launches rc=$?"
timeout 300 python tools/profile_kernels.py > $OUT/r02_profile_order.txt 2>&1; echo "profile_kernels rc=$?"; cat $OUT/r02_profile_order.txt | tail -20
REP=/tmp/prof_r02_all
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ps_ -o $REP python tools/profile_kernels.py > $OUT/r02_ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i $REP.ncu-rep --page raw --csv > $OUT/prof_r02_all_raw.csv 2>/dev/null
ncu -i $REP.ncu-rep --page source --csv -k regex:ps_apply_ldg -c 1 > $OUT/prof_r02_apply_source.csv 2>/dev/null
ncu -i $REP.ncu-rep --page source --csv -k regex:ps_step_switch -c 1 > $OUT/prof_r02_switch_source.csv 2>/dev/null
ls -la $OUT | tail -12
