#!/bin/bash
# SASS opcode evidence per kernel of libb200gate.so: tcgen05 (UTC*MMA), TMEM (LDTM/STTM), TMA (UTMALDG/UBLKCP),
# legacy tensor path (HMMA).  usage: tools/sass_histogram.sh > profiles/rNN_sass_opcodes.txt
set -e
LIB=${1:-vgate_b200/libb200gate.so}
echo "# $(basename $LIB): SASS opcode counts per kernel (cuobjdump -sass, sm_100a)"
cuobjdump -sass "$LIB" | awk '
  /Function :/ { fn=$3 }
  { if (match($0, /UTC[A-Z]*MMA|LDTM|STTM|UTMALDG|UTMASTG|UBLKCP|UTCBAR|HMMA|SYNCS|LDGSTS|UTMAPF|UTCCP|REDUX|CCTL/)) { op=substr($0, RSTART, RLENGTH); c[fn" "op]++ } }
  END { for (k in c) print k, c[k] }' | sort | while read fn op n; do
    printf "%-110s %-10s %s\n" "$(echo $fn | c++filt | cut -c1-110)" "$op" "$n"; done
