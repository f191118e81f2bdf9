#!/bin/bash
# profiles/sass_summary.txt: per kernel of libpn_b200.so, counts of the SASS mnemonics that prove the Blackwell paths
# (UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG/UTMASTG = TMA bulk tensor load/store, UTCBAR = commit)
SO=${1:-pacednegatives_b200/libpn_b200.so}
echo "# cuobjdump -sass $SO (sm_100a), $(date -u +%Y-%m-%d), commit $(git rev-parse --short HEAD 2>/dev/null)"
echo "# kernel | instructions | UTC*MMA | LDTM | STTM | UTMALDG | UTMASTG | UTCBAR | MUFU.EX2 | FFMA2/FADD2/FMUL2"
cuobjdump -sass "$SO" | awk '
  /Function :/ { if (name != "") out(); name=$3; n=mma=ldtm=sttm=ldg=stg=bar=ex2=f2=0; next }
  $1 ~ /^\/\*[0-9a-f][0-9a-f][0-9a-f][0-9a-f]\*\/$/ { n++
    if ($0 ~ /UTC[A-Z]*MMA/) mma++
    if ($0 ~ /LDTM/) ldtm++
    if ($0 ~ /STTM/) sttm++
    if ($0 ~ /UTMALDG/) ldg++
    if ($0 ~ /UTMASTG/) stg++
    if ($0 ~ /UTCBAR/) bar++
    if ($0 ~ /MUFU\.EX2/) ex2++
    if ($0 ~ /FFMA2|FADD2|FMUL2/) f2++ }
  function out() { printf "%s | %d | %d | %d | %d | %d | %d | %d | %d | %d\n", name, n, mma, ldtm, sttm, ldg, stg, bar, ex2, f2 }
  END { if (name != "") out() }' | c++filt | sort -t'|' -k3 -n -r
