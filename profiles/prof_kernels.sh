#!/bin/bash
# ncu --set full captures of the kernels named on the command line (one_kernel.py workloads), each after
# its plain run exited 0.  Usage: bash profiles/prof_kernels.sh TAG workload:kernel_regex[:extra args] ...
TAG=$1; shift
OUT=gpurun_out; mkdir -p $OUT
for spec in "$@"; do
  IFS=: read -r wl rx extra <<< "$spec"
  CMD="python profiles/one_kernel.py $wl reps=2 $extra"
  timeout 300 $CMD > $OUT/${TAG}_${wl}_plain.log 2>&1 || { echo "$wl plain run failed"; tail -3 $OUT/${TAG}_${wl}_plain.log; continue; }
  cat $OUT/${TAG}_${wl}_plain.log
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$rx -s 3 -c 1 -f \
      -o $OUT/${TAG}_${wl} $CMD > $OUT/${TAG}_${wl}_ncu.log 2>&1
  echo "$wl ncu exit $?"
done
