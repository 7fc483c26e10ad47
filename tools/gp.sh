#!/bin/sh
# retry wrapper around gpurun: exit code 3 = no slot right now (nothing charged) -> wait and retry
i=0
while [ $i -lt 20 ]; do
  /usr/local/graft/bin/gpurun "$@"
  rc=$?
  [ $rc -ne 3 ] && exit $rc
  i=$((i+1)); sleep 90
done
exit 3
