V="$1"; shift
for v in $V; do
  [ "$v" = "-" ] && v=""
  echo "== libvrt$v $@"
  VRT_LIB=$PWD/voxel-raytracing_b200/libvrt$v.so timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --min-timed-ms 300 "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']
print('ms/frame', round(d['ms_per_step'],3), {k: round(v,3) for k,v in r['stages_ms'].items()}, 'fields GiB', round(d['scene']['field_bytes']/2**30,3), 'load ms', round(d['scene']['load_ms']))"
done
