#!/bin/bash
# usage: grun.sh <tag> <timeout> <command...>   (retries on transient "no slot" answers)
tag=$1; shift; to=$1; shift
mkdir -p /root/repo/gpurun_out
for i in 1 2 3 4 5 6 7 8 9 10 11 12; do
  /usr/local/graft/bin/gpurun --gpus ${GPUS:-1} --timeout $to -- "$@" > /root/repo/gpurun_out/${tag}_call.log 2>&1
  rc=$?
  if [ $rc -ne 3 ] && ! grep -q "status=transient" /root/repo/gpurun_out/${tag}_call.log; then break; fi
  sleep 90
done
echo "rc=$rc" >> /root/repo/gpurun_out/${tag}_call.log
