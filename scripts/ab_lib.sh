# usage: scripts/ab_lib.sh OTHER_LIB.so [batches...] -> interleaved A/B of the in-tree library against another build
# of the same ABI (SC_B200_LIB), ms per step at each batch size
other=$1; shift
for bs in ${@:-2048 4096 8192}; do for x in new old new old; do
  if [ $x = old ]; then export SC_B200_LIB=$other; else unset SC_B200_LIB; fi
  echo -n "B=$bs $x: "; python bench.py --quick --steps 200 --warmup 20 --batch $bs 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['ms_per_step'],4))"
done; done
