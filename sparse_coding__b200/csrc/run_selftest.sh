#!/bin/bash
# Runs the GEMM-core self-test matrix, one process per configuration (a trapped kernel must not
# take the rest down).  Usage: run_selftest.sh [out_file]
BIN="$(dirname "$0")/bin/sc_selftest"
OUT="${1:-/dev/stdout}"
: > "$OUT"
run() { timeout 120 "$BIN" "$@" >> "$OUT" 2>&1 || echo "  -> exit $? for: $*" >> "$OUT"; }
#    M    N    K   G a_mn b_mn BN nseg
run 128  256   32  1  0 0 256 1
run 128  256  512  1  0 0 256 1
run 256  512  512  2  0 0 256 1
run 300  520  100  3  0 0 256 1
run 300  520  100  3  0 0 128 1
run 300  520  100  3  0 0  64 1
run 256  512  256  1  0 0 256 3
run 128  256   32  1  1 0 256 1
run 128  256   32  1  0 1 256 1
run 128  256   32  1  1 1 256 1
run 300  520  100  3  1 1 256 1
run 300  520  100  3  1 1 128 2
run 300  520  100  3  0 1  64 1
run 4096 4096 512  8  0 0 256 1
run 4096 4096 512  8  0 0 128 1
run 4096 512 4096  8  0 0 256 1
run 4096 512 4096  8  1 1 256 1
run 4096 4096 512  8  0 1 256 1
