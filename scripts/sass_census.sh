#!/bin/bash
# Per-kernel opcode census of the built library (no GPU needed): counts of the Blackwell-native instructions
# (UTC*MMA = tcgen05.mma, UTMALDG = TMA load, LDTM / STTM = tcgen05.ld / st) and of the legacy tensor path (HMMA) and
# MUFU per kernel, from `cuobjdump -sass`.  Usage: scripts/sass_census.sh [lib.so] > profiles/sass_census.txt
LIB=${1:-2handedafforder_b200/libb200sam.so}
echo "# cuobjdump -sass $LIB  ($(date -u +%Y-%m-%dT%H:%MZ), $(nvcc --version | tail -2 | head -1))"
printf "%-72s %8s %6s %8s %6s %6s %6s %6s %8s\n" kernel UTCxMMA .2CTA UTMALDG LDTM STTM HMMA MUFU instrs
cuobjdump -sass "$LIB" | (command -v c++filt > /dev/null && c++filt || cat) | awk '
function flush() {
  if (name != "") printf "%-72s %8d %6d %8d %6d %6d %6d %6d %8d\n", substr(name, 1, 72), mma, cta2, tma, ldtm, sttm, hmma, mufu, n
  tm += mma; t2 += cta2; tt += tma; tl += ldtm; ts += sttm; th += hmma; tu += mufu; tn += n
  mma = cta2 = tma = ldtm = sttm = hmma = mufu = n = 0
}
/Function :/ { flush(); name = $0; sub(/.*Function : /, "", name); gsub(/\(anonymous namespace\)::/, "", name)
               sub(/\(.*/, "", name); sub(/^void /, "", name) }
/^ +\/\*[0-9a-f]+\*\/ / {
  n++
  if ($0 ~ /UTC[A-Z]*MMA/) { mma++; if ($0 ~ /2CTA/) cta2++ }
  if ($0 ~ /UTMALDG/) tma++
  if ($0 ~ /LDTM/) ldtm++
  if ($0 ~ /STTM/) sttm++
  if ($0 ~ / HMMA/) hmma++
  if ($0 ~ /MUFU/) mufu++
}
END { flush(); printf "%-72s %8d %6d %8d %6d %6d %6d %6d %8d\n", "TOTAL", tm, t2, tt, tl, ts, th, tu, tn }'
