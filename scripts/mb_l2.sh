B=sparse_coding__b200/csrc/bin/sc_selftest
export SC_SELFTEST_NULL_EPI=1
for cfg in "4096 2048 512 8 0 0 256 1 2 2" "4096 2048 512 8 0 0 128 1 2 2" "2048 512 4096 8 0 0 256 1 2 2" "2048 512 4096 8 0 0 128 1 2 2" "4096 2048 512 1 0 0 256 1 2 2" "8192 8192 512 1 0 0 256 1 2 2"; do
  timeout 120 $B $cfg 2>&1 | sed 's/max_err.*bad=[0-9]*//'
done
