#!/bin/bash
# tools/gr.sh GPUS LOG TIMEOUT 'command' -- gpurun with retries on "busy" (exit code 3)
g=$1; log=$2; to=$3; shift 3
for i in $(seq 1 20); do
  if [ "$g" = 1 ]; then /usr/local/graft/bin/gpurun --timeout $to -- "$@" > $log 2>&1; else /usr/local/graft/bin/gpurun --gpus $g --timeout $to -- "$@" > $log 2>&1; fi
  rc=$?
  if [ $rc -ne 3 ]; then echo "rc=$rc" >> $log; exit $rc; fi
  sleep 60
done
echo "rc=3 (gave up)" >> $log; exit 3
