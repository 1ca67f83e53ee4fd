#!/bin/bash
# Runs every test of the given pytest file(s) in its own process (a trapped kernel
# kills the CUDA context, so one bad case must not take the others down).
# usage: scripts/run_isolated.sh <log> <pytest args...>
LOG=$1; shift
mkdir -p "$(dirname "$LOG")"
: > "$LOG"
IDS=$(python -m pytest "$@" --collect-only -q 2>/dev/null | grep '::')
for id in $IDS; do
  timeout 180 python -m pytest "$id" -q -x --no-header -p no:cacheprovider > /tmp/iso_one.log 2>&1
  rc=$?
  echo "rc=$rc $id" >> "$LOG"
  if [ $rc -ne 0 ]; then
    echo "---- $id" >> "$LOG.fail"
    tail -40 /tmp/iso_one.log >> "$LOG.fail"
  fi
done
echo "passed: $(grep -c '^rc=0' "$LOG") failed: $(grep -vc '^rc=0' "$LOG")"
