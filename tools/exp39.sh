B="python bench.py --no-cpu-baseline --steps 8 --warmup 3"
get() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(round(d["ms_per_step"],3), d["roofline"].get("stages_ms"))
PY
}
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
echo "== full"; $B > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; echo "== shard 0/8"; $B --shard 0 8 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json
echo "== cfg3"; $B --config cfg3 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json
echo "== cfg2"; $B --config cfg2 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json
