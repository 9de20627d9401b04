#!/bin/bash
# Profiling recipe (run under gpurun): plain run first, then ncu on the SAME command line.
#   tools/gpu_profile.sh <tag> [kernel-regex ...]
set -u
TAG=${1:-prof}; shift || true
mkdir -p gpurun_out
CMD="python bench.py --tran-steps 200 --steps 2 --warmup 1 --instances-per-gpu 2048 --no-cpu-baseline --no-e2e --stamp-instances 262144 --strong-total 0"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/${TAG}_plain.log; exit 1; }
tail -1 gpurun_out/${TAG}_plain.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launch.log 2>&1
echo "launch list rc=$?"
for K in "$@"; do
  ncu --set full --clock-control none --import-source on -k regex:$K -s 1 -c 1 -f -o gpurun_out/${TAG}_$K $CMD > gpurun_out/${TAG}_ncu_$K.log 2>&1
  echo "$K full rc=$?"
done
ls -la gpurun_out
