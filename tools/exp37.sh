B="python bench.py --no-cpu-baseline --steps 8 --warmup 3"
get() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(round(d["ms_per_step"],3), d["roofline"].get("stages_ms"))
PY
}
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for rep in 1 2; do echo "== full"; $B > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; echo "== shard 0/8"; $B --shard 0 8 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; done
echo "== shard 0/8 no overlap"; VRT_OVERLAP=0 $B --shard 0 8 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json
echo "== shard 0/8 late cap 16"; VRT_BOUNCE_CAP_LATE=16 $B --shard 0 8 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json
echo "== full late cap 16"; VRT_BOUNCE_CAP_LATE=16 $B > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json
echo "== full chain 1"; VRT_BOUNCE_CHAIN=1 $B > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json
echo "== shard chain 1"; VRT_BOUNCE_CHAIN=1 $B --shard 0 8 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json
