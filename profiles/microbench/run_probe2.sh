#!/bin/bash
cd "$(dirname "$0")"
for i in $(seq 18 47); do timeout 60 ./tcgen05_probe t3 $i | grep -v "^device\|probe done"; done
