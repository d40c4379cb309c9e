# marginal in-situ cost of each idempotent kernel of the fp16 step: step time with that kernel launched twice
for k in -1 0 1 2 3 -1; do
  echo -n "dup=$k "; SC_B200_DUP=$k python bench.py --steps 300 --warmup 20 --quick 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['ms_per_step'],4))"
done
