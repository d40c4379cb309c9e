#!/bin/bash
# Per-kernel counts of the SASS instruction classes that prove the Blackwell-native path
# (tcgen05.mma -> UTC*MMA, tcgen05.ld -> LDTM, TMA -> UTMALDG / UTMASTG / UTMAREDG, tcgen05.commit -> UTCBAR).
#   scripts/sass_summary.sh > profiles/r02_sass_summary.txt
SO="$(dirname "$0")/../sparse_coding__b200/libsc_b200.so"
cuobjdump -sass "$SO" | awk '
/Function :/ { name=$3; next }
/UTC[A-Z]*MMA/ { mma[name]++ }
/LDTM/ { ldtm[name]++ }
/UTMALDG/ { ldg[name]++ }
/UTMASTG/ { stg[name]++ }
/UTMAREDG/ { red[name]++ }
/UTCBAR/ { bar[name]++ }
/F2FP.SATFINITE/ { sat[name]++ }
END {
  printf "%-7s %-5s %-8s %-8s %-9s %-7s %-8s %s\n", "UTC*MMA", "LDTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UTCBAR", "F2FP.SAT", "kernel";
  for (k in mma) printf "%-7d %-5d %-8d %-8d %-9d %-7d %-8d %s\n", mma[k], ldtm[k], ldg[k], stg[k], red[k], bar[k], sat[k], k;
}' | (read -r header; echo "$header"; sort -k8 | c++filt | sed 's/sc::Gemm[A-Za-z]*, //g; s/(sc::GemmMaps.*//')
