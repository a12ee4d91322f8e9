#!/bin/bash
# gpurun with retries on "no box right now" (exit 3).  usage: tools/grun.sh <timeout> '<command>' [gpus]
T=$1; CMD=$2; G=${3:-1}
for i in 1 2 3 4 5 6 7 8 9 10 11 12; do
  if [ "$G" = "1" ]; then /usr/local/graft/bin/gpurun --timeout $T -- "$CMD"; else /usr/local/graft/bin/gpurun --gpus $G --timeout $T -- "$CMD"; fi
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 120
done
exit 3
