# usage: scripts/ab_env.sh VAR A B [batches...]  -> interleaved A/B of one environment switch (two values), ms per step
v=$1; a=$2; b=$3; shift 3
for bs in ${@:-2048 4096 8192}; do for x in $a $b $a $b; do
  echo -n "B=$bs $v=$x: "; env $v=$x python bench.py --quick --steps 200 --warmup 20 --batch $bs 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['ms_per_step'],4))"
done; done
