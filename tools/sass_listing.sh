#!/bin/bash
# Full SASS listing of every kernel in libb200sys.so (gzip) + the per-kernel mnemonic table.
#   usage: tools/sass_listing.sh [outdir]     (default profiles/r2/sass)
set -e
cd "$(dirname "$0")/.."
LIB=cs336_a2_systems_b200/lib/libb200sys.so
OUT=${1:-profiles/r2/sass}
mkdir -p "$OUT"
cuobjdump -sass "$LIB" | gzip -9 > "$OUT/libb200sys.sass.gz"
zcat "$OUT/libb200sys.sass.gz" | awk '/Function :/ {n++; name=$3} /^ +\/\*[0-9a-f]+\*\// {c[name]++} END {for (k in c) printf("%7d  %s\n", c[k], k)}' \
  | sort -k2 | c++filt > "$OUT/kernels_and_instruction_counts.txt"
bash tools/sass_evidence.sh "$LIB" > "$OUT/../sass_evidence.md"
echo "kernels: $(wc -l < "$OUT/kernels_and_instruction_counts.txt"), listing: $(du -h "$OUT/libb200sys.sass.gz" | cut -f1)"
