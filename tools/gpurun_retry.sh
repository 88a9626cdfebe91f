#!/bin/bash
# usage: tools/gpurun_retry.sh [--gpus N] <timeout> <script>   -- retries while the pod answers "transient" (nothing charged)
gp=""
if [ "$1" == "--gpus" ]; then gp="--gpus $2"; shift 2; fi
to=$1; script=$2
for attempt in 1 2 3 4 5 6 7 8 9 10 11 12; do
  out=$(gpurun $gp --timeout $to -- "bash $script" 2>&1)
  echo "$out" | tail -45
  if echo "$out" | grep -q "status=transient"; then echo "[retry $attempt] pod busy, sleeping"; sleep 150; continue; fi
  break
done
