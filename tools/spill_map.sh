#!/bin/bash
# usage: tools/spill_map.sh <mangled-kernel-substring>   (run from signalizer_b200/csrc)
# Lists local-memory (spill) instructions of one kernel by CUDA source line.
set -e
nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -cubin -o /tmp/sgz.cubin sgz_spectrum.cu 2>/dev/null
nvdisasm -g /tmp/sgz.cubin 2>/dev/null | awk -v k="$1" '/\.text\._ZN3sgz/{on = ($0 ~ k)} on' | grep -E "## File|STL|LDL" \
 | awk '/## File/{f=$0; next} {print f " || " $0}' \
 | sed -E 's/.*File "([^"]*)", line ([0-9]+).*\|\| *(\/\*[0-9a-f]+\*\/)? *([A-Z]+)[^;]*;.*/\1:\2 \4/' \
 | sed 's#.*/csrc/##; s#.*/include/##' | sort | uniq -c | sort -k2
