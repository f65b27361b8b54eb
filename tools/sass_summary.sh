#!/bin/sh
# Opcode histogram per kernel of libssb200.so (cuobjdump -sass): the tcgen05 / TMA / cluster instructions each kernel
# really contains.  Writes profiles/sass_summary.txt.   sh tools/sass_summary.sh
cd "$(dirname "$0")/.." || exit 1
SO=speech-separation_b200/ssb200/libssb200.so
cuobjdump -sass "$SO" | awk '
  /Function : / { fn=$3; next }
  /^ +\/\*[0-9a-f]+\*\/ +[A-Z@!]/ {
    op=$2; if (op ~ /^@/) op=$3; sub(/;$/, "", op); split(op, a, "."); base=a[1];
    if (base ~ /^(UTCHMMA|UTCQMMA|LDTM|STTM|UTMALDG|UTMASTG|UBLKCP|UTCBAR|SYNCS|STAS|UTCCP|LDGSTS|HMMA|FFMA2|FFMA|MUFU|SHFL|ATOM|RED|LDG|STG|LDS|STS|BAR|MEMBAR|FENCE|ACQBULK|UCGABAR|CCTL|ERRBAR)$/) cnt[fn" "base]++
    tot[fn]++
  }
  END { for (k in cnt) { split(k, b, " "); printf "%s %s %d\n", b[1], b[2], cnt[k] } for (f in tot) printf "%s TOTAL %d\n", f, tot[f] }' \
  | sort | c++filt | awk '{ key=$0; sub(/ [A-Z0-9]+ [0-9]+$/, "", key); n=split($0, p, " "); line[key]=line[key] " " p[n-1] "=" p[n] } END { for (k in line) print k ":" line[k] }' | sort > profiles/sass_summary.txt
wc -l profiles/sass_summary.txt
