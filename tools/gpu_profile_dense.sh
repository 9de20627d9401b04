#!/bin/bash
# Profiling recipe of the large-matrix path (run under gpurun): plain run first, then ncu on the SAME command line.
#   tools/gpu_profile_dense.sh <tag>
set -u
TAG=${1:-dense}
mkdir -p gpurun_out
CMD="python tools/run_config5.py --instances 592 --steps 12 --cpu-steps 0"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/${TAG}_plain.log; exit 1; }
tail -1 gpurun_out/${TAG}_plain.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 50 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launch.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:tran_dense_kernel -c 1 -f -o gpurun_out/${TAG}_tran_dense $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "full rc=$?"
[ -x build/dn_prof ] && build/dn_prof 404 296 | tee gpurun_out/${TAG}_dn_prof.txt
ls -la gpurun_out
