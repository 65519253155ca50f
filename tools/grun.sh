#!/bin/bash
# usage: tools/grun.sh TIMEOUT [--gpus N] -- 'command'   : retries while the pod answers busy / transient
T=$1; shift
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun --timeout $T "$@" > /tmp/grun_last.txt 2>&1
  if grep -q "nothing was charged" /tmp/grun_last.txt; then sleep 90; continue; fi
  break
done
cat /tmp/grun_last.txt
