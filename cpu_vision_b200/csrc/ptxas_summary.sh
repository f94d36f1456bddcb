#!/bin/sh
# usage: tools_ptxas.sh <file.ptxas.log>  -> "kernel regs smem" per entry
awk '/Compiling entry function/ {match($0, /'"'"'[^'"'"']+'"'"'/); name=substr($0, RSTART+1, RLENGTH-2)} /Used [0-9]+ registers/ {print name, $0}' "$1" | sed 's/ptxas info    : //' | while read n rest; do echo "$(echo $n | c++filt | cut -c1-90) | $rest"; done
