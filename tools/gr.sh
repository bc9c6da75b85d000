#!/bin/bash
# usage: tools/gr.sh <logfile> <gpurun args...> — runs gpurun, retrying while it answers "busy / no box" (exit 3) or a transient refusal
LOG=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" > "$LOG" 2>&1; rc=$?
  if [ $rc -eq 3 ] || grep -q "status=transient" "$LOG"; then sleep 45; continue; fi
  break
done
tail -60 "$LOG"; exit $rc
