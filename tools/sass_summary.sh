#!/bin/bash
# Per shipped kernel: counts of the Blackwell instructions in the SASS of libb200attn.so (cuobjdump -sass), the evidence
# that the hot path runs on tcgen05 / TMEM / TMA:  UTCHMMA = tcgen05.mma (UTCQMMA would be the fp8/fp4 kinds),
# LDTM / STTM = tcgen05.ld / st, UTMALDG / UTMASTG = TMA tensor load / store, UBLKCP = bulk copy, UBLKRED / UTMAREDG = bulk
# reduce, UTCBAR = tcgen05.commit, SYNCS = mbarrier ops.
#   bash tools/sass_summary.sh > profiles/r2_sass_summary.txt
cd "$(dirname "$0")/.."
LIB=${1:-cuda_flashattention_b200/libb200attn.so}
echo "# $LIB  ($(stat -c %s $LIB) bytes)  $(date -u +%F)"
cuobjdump -sass $LIB | awk '
/Function :/ { fn=$3; order[++n]=fn }
{ for (i=1;i<=NF;i++) { t=$i; sub(/\..*/,"",t);
    if (t ~ /^(UTCHMMA|UTCQMMA|UTCOMMA|LDTM|STTM|UTMALDG|UTMASTG|UBLKCP|UBLKRED|UTMAREDG|UTCBAR|SYNCS|MUFU|HMMA|FFMA2|FFMA)$/) c[fn,t]++ } }
END { printf "%-10s %-8s %-6s %-6s %-8s %-8s %-7s %-8s %-7s %-6s %-6s %-6s %-6s  %s\n","UTCHMMA","LDTM","STTM","UTMALDG","UTMASTG","UBLKCP","UBLKRED","UTMAREDG","UTCBAR","SYNCS","MUFU","HMMA","FFMA2","kernel";
  for (k=1;k<=n;k++) { fn=order[k];
    printf "%-10d %-8d %-6d %-6d %-8d %-8d %-7d %-8d %-7d %-6d %-6d %-6d %-6d  %s\n", c[fn,"UTCHMMA"]+c[fn,"UTCQMMA"]+c[fn,"UTCOMMA"],c[fn,"LDTM"],c[fn,"STTM"],c[fn,"UTMALDG"],c[fn,"UTMASTG"],c[fn,"UBLKCP"],c[fn,"UBLKRED"],c[fn,"UTMAREDG"],c[fn,"UTCBAR"],c[fn,"SYNCS"],c[fn,"MUFU"],c[fn,"HMMA"],c[fn,"FFMA2"], fn } }' | while read -r line; do
  fn=$(echo "$line" | awk '{print $NF}'); case "$fn" in _Z*) echo "$(echo "$line" | sed "s|$fn||")$(echo $fn | c++filt | cut -c1-110)";; *) echo "$line";; esac; done
