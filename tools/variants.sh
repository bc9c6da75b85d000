# usage: tools/variants.sh "<suffix> <suffix> ..." [bench args]  (GPU box): one bench line per build variant voxel-raytracing_b200/libvrt<suffix>.so
# ("" = the product library). Variants are built here with `make OUT=libvrt_x.so BUILD=build_x NVEXTRA=-D...` and loaded through VRT_LIB.
V="$1"; shift
for v in $V; do
  [ "$v" = "-" ] && v=""
  echo "== libvrt$v $@"
  VRT_LIB=$PWD/voxel-raytracing_b200/libvrt$v.so timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --min-timed-ms 300 "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']
print('ms/frame', round(d['ms_per_step'],3), 'latency', round(d['latency_ms_single_frame'],3), 'e2e', round(d['e2e']['ms_per_step'],3), {k: round(v,3) for k,v in r['stages_ms'].items()})"
done
