#!/bin/bash
# SASS evidence of the built library (run in the build container): per kernel, counts of the bulk-copy (TMA engine),
# mbarrier, named-barrier and local-memory instructions.  Output -> profiles/r2_sass_evidence.txt
LIB=${1:-omniisaacgymenvs_b200/liboige_b200.so}
echo "# cuobjdump -sass $LIB  (nvcc $(nvcc --version | grep release | sed 's/.*release //'))"
cuobjdump -lelf "$LIB" | sed 's/^/# /'
cuobjdump -sass "$LIB" | awk '
/Function :/ { name=$3; sub(/_ZN[0-9a-zA-Z_]*oige[0-9]*/, "", name); fn=substr($3,1,150); next }
fn != "" {
  if ($0 ~ /UBLKCP/) ublk[fn]++
  if ($0 ~ /SYNCS/) syncs[fn]++
  if ($0 ~ /BAR\.(SYNC|ARV)/) bar[fn]++
  if ($0 ~ /LDL|STL/) loc[fn]++
  if ($0 ~ /^ +\/\*[0-9a-f]+\*\/ /) n[fn]++
}
END { printf "%8s %7s %6s %5s %7s  kernel\n", "sass", "UBLKCP", "SYNCS", "BAR", "LDL/STL"; for (f in n) printf "%8d %7d %6d %5d %7d  %s\n", n[f], ublk[f], syncs[f], bar[f], loc[f], f }' | sort -k6
