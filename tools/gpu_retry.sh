#!/bin/bash
# tools/gpu_retry.sh OUTFILE TIMEOUT [--gpus N] -- CMD : retry gpurun while the pod answers "transient" (busy), output to OUTFILE
OUT=$1; shift; TMO=$1; shift
EXTRA=""
if [ "$1" == "--gpus" ]; then EXTRA="--gpus $2"; shift; shift; fi
shift
for k in $(seq 1 40); do
  /usr/local/graft/bin/gpurun $EXTRA --timeout $TMO -- "$@" > "$OUT" 2>&1
  if ! grep -q "status=transient" "$OUT"; then exit 0; fi
  sleep 45
done
