#!/bin/bash
# tools/grun.sh with retries while the pod answers "transient" (nothing charged): tools/grun_retry.sh LOG [gpurun flags] -- 'cmd'
log="$1"; shift
for i in 1 2 3 4 5 6 7 8; do
  "$(dirname "$0")/grun.sh" "$@" > "$log" 2>&1
  if ! grep -q "status=transient" "$log"; then exit 0; fi
  sleep 90
done
