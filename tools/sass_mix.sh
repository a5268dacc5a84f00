#!/bin/bash
# Opcode histogram of a kernel in libmtsb200.so (or $2): tools/sass_mix.sh <mangled-name-substring> [lib]
LIB=${2:-/root/repo/mitsuba_b200/libmtsb200.so}
FN=$(cuobjdump -sass $LIB | grep "Function :" | grep "$1" | head -1 | awk '{print $3}')
echo "function $FN"
cuobjdump -sass -fun "$FN" $LIB | grep -E "^\s+/\*[0-9a-f]{4}\*/" | sed 's/\/\* 0x[0-9a-f]* \*\///' > /tmp/sass_$$.txt
echo "total instructions: $(wc -l < /tmp/sass_$$.txt)"
awk '{op=$2; if (op ~ /^@/) op=$3; sub(/\..*/,"",op); print op}' /tmp/sass_$$.txt | sort | uniq -c | sort -rn | head -${3:-30}
cp /tmp/sass_$$.txt /tmp/sass_last.txt; rm /tmp/sass_$$.txt
