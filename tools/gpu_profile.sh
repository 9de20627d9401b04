#!/bin/bash
# Round-1 profiling recipe (run under gpurun): plain run first, then ncu on the SAME command line.
set -u
mkdir -p gpurun_out
CMD="python bench.py --tran-steps 200 --steps 2 --warmup 1 --instances-per-gpu 2048 --no-cpu-baseline --no-e2e --stamp-instances 262144"
$CMD > gpurun_out/prof_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/prof_plain.log; exit 1; }
tail -1 gpurun_out/prof_plain.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:tran_kernel -s 1 -c 1 -o gpurun_out/prof_tran $CMD > gpurun_out/ncu_tran.log 2>&1
echo "tran full rc=$?"
ncu --set full --clock-control none --import-source on -k regex:stamp_csr -s 4 -c 1 -o gpurun_out/prof_stamp $CMD > gpurun_out/ncu_stamp.log 2>&1
echo "stamp full rc=$?"
ls -la gpurun_out
