# usage: tools/env_exp.sh "VAR=val VAR=val ..." [bench args]: one bench line per environment setting ("-" = none)
E="$1"; shift
for e in $E; do
  [ "$e" = "-" ] && e="A=1"
  echo "== $e $@"
  env $e timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --min-timed-ms 300 "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']
print('ms/frame', round(d['ms_per_step'],3), 'latency', round(d['latency_ms_single_frame'],3), {k: round(v,3) for k,v in r['stages_ms'].items()})"
done
