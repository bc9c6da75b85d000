B="python bench.py --no-cpu-baseline --steps 8 --warmup 3"
get() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(round(d["ms_per_step"],3), d["roofline"].get("stages_ms"))
PY
}
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for c in 4 3 2 1 0; do echo "== full chain $c"; VRT_BOUNCE_CHAIN=$c $B > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; done
for c in 4 3 2 1 0; do echo "== shard 0/8 chain $c"; VRT_BOUNCE_CHAIN=$c $B --shard 0 8 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; done
for c in 4 2 1; do echo "== shard 0/8 no overlap chain $c"; VRT_OVERLAP=0 VRT_BOUNCE_CHAIN=$c $B --shard 0 8 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; done
