#!/bin/bash
# usage: tools/gpurun_retry.sh <timeout> <logfile> <command...> : retries while the pod answers busy / transient (nothing charged)
T=$1; LOG=$2; shift 2
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $T -- "$@" > $LOG 2>&1
  if grep -q "status=transient\|rc=3\|no box\|busy" $LOG && ! grep -q "charged=[1-9]" $LOG; then sleep 90; continue; fi
  break
done
echo "gpurun_retry finished (attempt $i)" >> $LOG
