# engine statistics (stats build) + side-stream quota sweep on a 1/8 shard and on the full frame
B="python bench.py --no-cpu-baseline --steps 5 --warmup 3"
get() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(round(d["ms_per_step"],3), d["roofline"].get("stages_ms"))
PY
}
echo "== stats, full frame, cfg4"; VRT_LIB=tools/libvrt_stats.so VRT_DEBUG_STATS=1 $B --steps 1 --warmup 3 > gpurun_out/x.json 2> gpurun_out/x.err; grep "engine stats" gpurun_out/x.err | tail -1
echo "== stats, cfg3"; VRT_LIB=tools/libvrt_stats.so VRT_DEBUG_STATS=1 $B --config cfg3 --steps 1 --warmup 3 > gpurun_out/x.json 2> gpurun_out/x.err; grep "engine stats" gpurun_out/x.err | tail -1
for q in 256 512 1024 2048 4096; do echo "== shard 0/8 quota $q"; VRT_ENGINE_QUOTA=$q $B --shard 0 8 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; done
echo "== shard 0/8 overlap off"; VRT_OVERLAP=0 $B --shard 0 8 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json
for q in 512 2048; do echo "== shard 0/4 quota $q"; VRT_ENGINE_QUOTA=$q $B --shard 0 4 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; done
