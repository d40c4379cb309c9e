# mainloop alone (null epilogue) on the encode shape (4096 x 2048 x 512, 8 groups, fp16, CTA pairs) with 256- and 128-wide
# tiles: the same flops with 1.5x the operand bytes L2 -> SM.  Also long-K shapes for comparison.
B=sparse_coding__b200/csrc/bin/sc_selftest
export SC_SELFTEST_NULL_EPI=1 SC_SELFTEST_REPS=50
for cfg in "4096 2048 512 8 0 0 256 1 2 2" "4096 2048 512 8 0 0 128 1 2 2" "2048 512 4096 8 0 0 256 1 2 2" "2048 512 4096 8 0 0 128 1 2 2" "8192 8192 8192 1 0 0 256 1 2 2" "8192 8192 8192 1 0 0 128 1 2 2"; do
  echo -n "$cfg : "; timeout 120 $B $cfg 2>&1 | sed 's/max_err.*bad=[0-9]*//' | tail -1
done
