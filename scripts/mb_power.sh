# Is the mainloop power-limited?  The same GEMM (null epilogue, 2000 back-to-back launches = 0.1 s) on random and on
# all-zero operands: identical instruction stream and memory traffic, only the data toggling differs.
B=sparse_coding__b200/csrc/bin/sc_selftest
export SC_SELFTEST_NULL_EPI=1 SC_SELFTEST_REPS=2000
for z in "" 1 "" 1; do
  echo -n "zero=${z:-0}: "; env ${z:+SC_SELFTEST_ZERO=1} timeout 200 $B 4096 2048 512 8 0 0 256 1 2 2 2>&1 | sed 's/max_err.*bad=[0-9]*//' | tail -1
done
