#!/bin/bash
# gpurun with retries while the pod answers "busy" (exit code 3): tools/gr.sh <log name> <timeout s> <command...>
name=$1; shift; to=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$to" "$@" > "gpurun_out/${name}_call.log" 2>&1
  rc=$?
  if [ $rc -ne 3 ] && ! grep -q "status=transient" "gpurun_out/${name}_call.log"; then break; fi
  sleep 45
done
echo "rc=$rc" >> "gpurun_out/${name}_call.log"
