#!/bin/bash
# Runs HERE (the build container): one gpurun call, retried while the pod answers "no box / slot free"
# (exit 3) or "transient".   tools/gpu_retry.sh <gpus> <log> <timeout-seconds> '<command>'
gpus=$1; log=$2; tmo=$3; shift 3
for try in $(seq 1 40); do
  if [ "$gpus" = "1" ]; then
    /usr/local/graft/bin/gpurun --timeout $tmo -- "$@" > $log 2>&1
  else
    /usr/local/graft/bin/gpurun --gpus $gpus --timeout $tmo -- "$@" > $log 2>&1
  fi
  rc=$?
  if [ $rc -ne 3 ] && ! grep -q "status=transient" $log; then echo "gpurun exit $rc (try $try)" >> $log; exit $rc; fi
  sleep 90
done
echo "gave up" >> $log
exit 3
