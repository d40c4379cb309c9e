B=sparse_coding__b200/csrc/bin/sc_selftest
for mode in 1 11 12 14 15 16 17 18; do
  echo -n "mode=$mode: "; SC_SELFTEST_H=2 SC_SELFTEST_NULL_EPI=$mode timeout 120 $B 4096 2048 512 8 0 0 256 1 2 2 2>&1 | sed 's/max_err.*bad=[0-9]*//' | tail -1
done
