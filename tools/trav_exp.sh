# usage: tools/trav_exp.sh "<traversal ids>" [bench args]: one bench line per traversal back-end (0 default, 5 linear octant fields, 6 two-level)
T="$1"; shift
for t in $T; do
  echo "== traversal $t $@"
  timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --min-timed-ms 300 --traversal $t "$@" 2>gpurun_out/trav_exp.err | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']
    print('ms/frame', round(d['ms_per_step'],3), 'latency', round(d['latency_ms_single_frame'],3), {k: round(v,3) for k,v in r['stages_ms'].items()}, 'scene GiB', round(d['scene_bytes']/2**30,2), 'fields GiB', round(d['scene']['field_bytes']/2**30,3), 'peak GiB', round(d['scene']['peak_build_bytes']/2**30,1), 'load ms', round(d['scene']['load_ms']))
except Exception as e: print('failed', e)
" || tail -3 gpurun_out/trav_exp.err
done
