B="python bench.py --no-cpu-baseline --steps 8 --warmup 3"
get() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(round(d["ms_per_step"],3), d["roofline"].get("stages_ms"))
PY
}
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for c in 32 64 128 512; do echo "== full late cap $c"; VRT_BOUNCE_CAP_LATE=$c $B > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; done
for c in 32 128; do echo "== shard 0/8 late cap $c"; VRT_BOUNCE_CAP_LATE=$c $B --shard 0 8 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; done
for c in 32 128; do echo "== shard 0/8 no overlap late cap $c"; VRT_OVERLAP=0 VRT_BOUNCE_CAP_LATE=$c $B --shard 0 8 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; done
