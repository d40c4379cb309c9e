B=sparse_coding__b200/csrc/bin/sc_selftest
for h in 2 4; do for mode in 10 11 12; do
  echo -n "H=$h mode=$mode: "; SC_SELFTEST_H=$h SC_SELFTEST_NULL_EPI=$mode timeout 120 $B 4096 2048 512 8 0 0 256 1 2 2 2>&1 | sed 's/max_err.*bad=[0-9]*//' | tail -1
done; done
for mode in 20 21; do
  echo -n "mode=$mode: "; SC_SELFTEST_NULL_EPI=$mode timeout 120 $B 4096 2048 512 8 0 0 256 1 2 2 2>&1 | sed 's/max_err.*bad=[0-9]*//' | tail -1
done
