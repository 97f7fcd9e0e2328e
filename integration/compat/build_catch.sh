#!/bin/bash
# Builds build/compat/catch_phylohmm: the reference's own TEST_CASE("PhyloHMM") (test/test.cpp:750-1399), compiled
# from where it lies against the product's classes through integration/compat/linearham_compat.hpp and linked with
# linearham_b200/liblinearham_b200.so. Needs the reference tree (Eigen and catch.hpp are taken from it too).
#   usage: integration/compat/build_catch.sh [reference_root]
set -e
REF=${1:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
ROOT=$(cd "$HERE/../.." && pwd)
OUT=$ROOT/build/compat
CXX=${CXX:-$( [ -x /opt/gcc/bin/g++ ] && echo /opt/gcc/bin/g++ || echo g++ )}
mkdir -p "$OUT"
start=$(grep -n '^TEST_CASE("PhyloHMM"' "$REF/test/test.cpp" | cut -d: -f1)
[ -n "$start" ] || { echo "TEST_CASE(\"PhyloHMM\") not found in $REF/test/test.cpp" >&2; exit 1; }
{ cat "$HERE/catch_prelude.hpp"; echo "#line $start \"$REF/test/test.cpp\""; tail -n +"$start" "$REF/test/test.cpp"; } > "$OUT/catch_phylohmm.cpp"
$CXX -O1 -std=c++17 -w -I"$HERE" -I"$ROOT/linearham_b200/csrc" -I"$REF/lib/eigen" -I"$REF/test" \
  "$OUT/catch_phylohmm.cpp" -o "$OUT/catch_phylohmm" -L"$ROOT/linearham_b200" -llinearham_b200 \
  -Wl,-rpath,'$ORIGIN/../../linearham_b200' -lpthread
rm -f "$OUT/catch_phylohmm.cpp"
echo "$OUT/catch_phylohmm"
