#!/bin/bash
# Per-kernel counts of the SASS mnemonics that prove the Blackwell datapath (tcgen05 -> UTC*MMA, tcgen05.ld -> LDTM,
# TMA -> UTMALDG/UTMASTG, multimem -> *.MULTIMEM / peer LDG/STG .SYS).  usage: tools/sass_evidence.sh > profiles/sass_evidence.md
LIB=${1:-cs336_a2_systems_b200/lib/libb200sys.so}
echo "# SASS evidence for $(basename $LIB) (cuobjdump -sass, sm_100a)"
echo
echo "| kernel | UTCHMMA | UTCBAR | LDTM | UTMALDG | SYNCS | REDG | LDG/STG .SYS | MULTIMEM |"
echo "|---|---|---|---|---|---|---|---|---|"
cuobjdump -sass "$LIB" 2>/dev/null | awk '
/Function :/ { if (name != "") print_row(); name=$3; for (k in c) delete c[k]; next }
{ if ($0 ~ /UTCHMMA/) c["a"]++; if ($0 ~ /UTCBAR/) c["b"]++; if ($0 ~ /LDTM/) c["c"]++; if ($0 ~ /UTMALDG/) c["d"]++;
  if ($0 ~ /SYNCS\./) c["e"]++; if ($0 ~ /REDG/) c["f"]++; if ($0 ~ /(LDG|STG)[^;]*\.SYS/) c["g"]++; if ($0 ~ /MULTIMEM|\.MMEM|LDGMC|REDGMC|STGMC/) c["h"]++ }
function print_row() { if (c["a"]+c["c"]+c["d"]+c["g"]+c["h"]+c["f"] > 0) printf("| `%s` | %d | %d | %d | %d | %d | %d | %d | %d |\n", name, c["a"], c["b"], c["c"], c["d"], c["e"], c["f"], c["g"], c["h"]) }
END { print_row() }' | sed 's/_ZN4b200//' | c++filt -_ 2>/dev/null
