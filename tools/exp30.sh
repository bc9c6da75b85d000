B="python bench.py --no-cpu-baseline --steps 5 --warmup 3"
get() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(round(d["ms_per_step"],3), d["roofline"].get("stages_ms"))
PY
}
for pad in 0 20 32 50 90; do for q in 1024 2048; do echo "== shard 0/8 pad $pad KB quota $q"; VRT_SIDE_PAD_KB=$pad VRT_ENGINE_QUOTA=$q $B --shard 0 8 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; done; done
for pad in 0 32 50; do echo "== shard 0/4 pad $pad"; VRT_SIDE_PAD_KB=$pad $B --shard 0 4 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; done
for pad in 32 50; do echo "== full overlap=1 pad $pad"; VRT_OVERLAP=1 VRT_SIDE_PAD_KB=$pad $B > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; done
