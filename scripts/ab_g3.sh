for cfg in "SC_B200_RESA=1 SC_B200_G3_SLOTS=3" "SC_B200_RESA=0 SC_B200_G3_SLOTS=3" "SC_B200_RESA=1 SC_B200_G3_SLOTS=4" "SC_B200_RESA=0 SC_B200_G3_SLOTS=4" "SC_B200_RESA=1 SC_B200_G3_SLOTS=3" "SC_B200_RESA=1 SC_B200_G3_SLOTS=4"; do
  echo -n "$cfg: "; env $cfg python bench.py --quick --steps 300 --warmup 20 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['ms_per_step'],4))"
done
