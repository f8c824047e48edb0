# A/B of one environment switch on the default bench: prints ms per step, the per-family eager kernel times and clocks.
#   bash scripts/ab_env.sh VAR "val1 val2 ..." [repeats] [extra bench args]
VAR=$1; VALS=$2; REP=${3:-2}; shift 3 2>/dev/null
for r in $(seq $REP); do for v in $VALS; do
  env $VAR=$v timeout -s KILL 200 python bench.py --steps 30 "$@" 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read())
k = d['kernel_ms_per_chunk']
print('$VAR=$v', 'ms/step %.3f' % d['ms_per_step'], 'e2e %.0f' % d['e2e']['value'], {a: round(b, 3) for a, b in k.items()}, 'sum %.3f' % sum(k.values()), d['clocks']['sm_mhz'], d['clocks']['reasons'])"
done; done
