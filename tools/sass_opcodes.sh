#!/bin/bash
# Opcode histogram of every kernel in the shipped library (evidence for what the sm_100a code uses:
# FFMA2 / FMUL2 / FADD2 packed fp32, REDG vector reductions, UBLKCP / SYNCS bulk copies + mbarriers, ...).
#   tools/sass_opcodes.sh > profiles/r02_sass_opcodes.txt
LIB=${1:-cpu_vision_b200/libb200det.so}
echo "# cuobjdump -sass $LIB: per kernel, SASS instructions by opcode (top 14) -- $(date -u +%F)"
cuobjdump -sass "$LIB" | awk '
/Function :/ { if (name != "") flush(); name=$3; delete c; n=0; next }
/^[[:space:]]+\/\*[0-9a-f]{4}\*\// { op=$2; if (op ~ /^@/) op=$3; sub(/\..*/, "", op); sub(/;/, "", op); c[op]++; n++ }
function flush(   k, line, m, best) {
  cmd="c++filt " name; cmd | getline dem; close(cmd)
  printf "%s\n  %d instructions:", substr(dem, 1, 150), n
  for (i = 0; i < 14; i++) { best=""; m=0; for (k in c) if (c[k] > m) { m=c[k]; best=k } if (best == "") break; printf " %s %d", best, m; delete c[best] }
  printf "\n"
}
END { if (name != "") flush() }'
echo "# totals over the library"
cuobjdump -sass "$LIB" | grep -E "^\s+/\*[0-9a-f]{4}\*/" | awk '{op=$2; if (op ~ /^@/) op=$3; sub(/\..*/, "", op); sub(/;/, "", op); c[op]++} END {for (k in c) print c[k], k}' | sort -rn | head -40
