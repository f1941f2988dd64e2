#!/bin/bash
# Per-kernel counts of the Blackwell-native SASS mnemonics in the built library (runs anywhere: cuobjdump only).
#   bash profiles/sass_extract.sh > profiles/r2_sass_extract.txt
LIB=${1:-l-msfc-v2-training_b200/libb200lmsfc.so}
echo "# cuobjdump -sass $LIB  ($(date -u +%F)): tcgen05 = UTCHMMA (MMA) / UTCBAR (commit) / LDTM (TMEM load); TMA = UTMALDG (load) / UTMAREDG (reduce-add)"
cuobjdump -sass "$LIB" | awk '
  /Function :/ { fn=$3; next }
  /UTCHMMA|UTCBAR|LDTM|UTMALDG|UTMAREDG|UTMASTG|SYNCS|UCGABAR|LDGSTS|REDG|MULTIMEM/ {
    m=$0; sub(/^[^A-Z]*/, "", m); split(m, a, /[ .;]/); key=a[1];
    if (match($0, /UTCHMMA|UTCBAR|LDTM|UTMALDG|UTMAREDG|UTMASTG|SYNCS|UCGABAR|LDGSTS|MULTIMEM/)) key=substr($0, RSTART, RLENGTH);
    cnt[fn "\t" key]++ }
  END { for (k in cnt) print k "\t" cnt[k] }' | sort | c++filt | awk -F'\t' '{printf "%-90s %-10s %s\n", substr($1,1,90), $2, $3}'
echo
echo "# first tcgen05.mma / TMA lines of igemm_fwd_kernel<256,false> and wgrad_kernel<2>"
cuobjdump -sass "$LIB" | awk '/Function :/ {fn=$3} /UTCHMMA|UTMALDG|UTMAREDG|LDTM/ { if (fn ~ /igemm_fwd_kernelILi256ELb0/ || fn ~ /wgrad_kernelILi2/) { c[fn]++; if (c[fn] <= 6) print fn ": " $0 } }' | c++filt
