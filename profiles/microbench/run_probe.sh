#!/bin/bash
# Runs every probe case in its own process (a faulting case must not poison the others).
cd "$(dirname "$0")"
out=${1:-/dev/stdout}
{
  timeout 120 ./tcgen05_probe t1
  timeout 60 ./tcgen05_probe t4
  for i in $(seq 0 17); do timeout 60 ./tcgen05_probe t3 $i | grep -v "^device\|probe done"; done
} > "$out" 2>&1
