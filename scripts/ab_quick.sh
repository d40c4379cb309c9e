# usage: scripts/ab_quick.sh VAR   -> A/B of one 0/1 switch at three batch sizes
v=$1
for b in 2048 4096 8192; do for x in 1 0 1 0; do
  echo -n "B=$b $v=$x: "; env $v=$x python bench.py --quick --steps 200 --warmup 20 --batch $b 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['ms_per_step'],4))"
done; done
