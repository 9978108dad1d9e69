#!/bin/bash
# usage: tools/gpurun_retry_n.sh <gpus> <timeout> <logfile> <command...> : like gpurun_retry.sh on N GPUs of one box
N=$1; T=$2; LOG=$3; shift 3
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --gpus $N --timeout $T -- "$@" > $LOG 2>&1
  if grep -q "status=transient\|rc=3\|no box\|busy" $LOG && ! grep -q "charged=[1-9]" $LOG; then sleep 90; continue; fi
  break
done
echo "gpurun_retry finished (attempt $i)" >> $LOG
