# A/B of the fp16 step's switches at three batch sizes (ms per step, CUDA events, 300 steps)
for b in 2048 4096 8192; do
  for cfg in "SC_B200_RESA=1 SC_B200_FOLD=1" "SC_B200_RESA=0 SC_B200_FOLD=1" "SC_B200_RESA=1 SC_B200_FOLD=0" "SC_B200_RESA=0 SC_B200_FOLD=0"; do
    echo -n "B=$b $cfg: "; env $cfg python bench.py --quick --steps 300 --warmup 20 --batch $b 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['ms_per_step'],4))"
  done
done
