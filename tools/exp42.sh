B="python bench.py --no-cpu-baseline --steps 12 --warmup 3"
get() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(round(d["ms_per_step"],3), "e2e", round(d["e2e"]["ms_per_step"],3), d["roofline"].get("stages_ms"))
PY
}
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for f in 3 4; do echo "== full F=$f"; $B --frames-in-flight $f > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; done
for f in 3 4; do echo "== shard 0/8 F=$f"; $B --shard 0 8 --frames-in-flight $f > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; done
for f in 3 4; do echo "== shard 0/4 F=$f"; $B --shard 0 4 --frames-in-flight $f > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; done
echo "== cfg5 F=3"; python bench.py --no-cpu-baseline --config cfg5 --steps 8 --warmup 3 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; tail -2 gpurun_out/x.err
