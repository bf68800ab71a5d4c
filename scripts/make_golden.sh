#!/bin/bash
# Copies the reference's own test FIXTURES (data files and *.expected.* golden
# outputs, not its sources or scripts) from /root/reference/fxtract_test into
# tests/golden/fxtract_test so the suite can run where /root/reference does not
# exist (the GPU box).  Also records, under tests/golden/ref_out/, outputs of the
# reference binary built by oracle/Makefile (oracle/_ref/fxtract) for a set of
# extra command lines that the reference's run.sh does not cover.
set -euo pipefail
REF=${REF:-/root/reference}
ROOT=$(cd "$(dirname "$0")/.." && pwd)
DST=$ROOT/tests/golden/fxtract_test
mkdir -p "$DST"
for f in "$REF"/fxtract_test/*.fa "$REF"/fxtract_test/*.fq "$REF"/fxtract_test/*.txt \
         "$REF"/fxtract_test/*.expected "$REF"/fxtract_test/*.expected.*; do
  [ -f "$f" ] && cp "$f" "$DST/"
done
mkdir -p "$ROOT/tests/golden/patterns"
for f in random_100_30bp_patterns.txt random_100_60bp_patterns.txt random_100_patterns_minus_small.txt \
         random_1000_30bp_patterns.txt random_1000_patterns_minus_small.txt; do
  cp "$REF/fxtract_test/speed_test/$f" "$ROOT/tests/golden/patterns/"
done
chmod -R u+w "$ROOT/tests/golden"
echo "fixtures copied to $DST"
