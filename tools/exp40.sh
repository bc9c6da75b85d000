B="python bench.py --no-cpu-baseline --steps 10 --warmup 3"
get() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(round(d["ms_per_step"],3), "e2e", round(d["e2e"]["ms_per_step"],3), d["roofline"].get("stages_ms"))
PY
}
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for rc in 62 30 100 150; do echo "== full roundcost $rc"; VRT_ENGINE_ROUNDCOST=$rc $B > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; done
echo "== full serial"; $B --serial-frames > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json
echo "== shard 0/8"; $B --shard 0 8 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json
echo "== shard 0/8 serial"; $B --shard 0 8 --serial-frames > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json
echo "== shard 0/8 no side overlap"; VRT_OVERLAP=0 $B --shard 0 8 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json
echo "== cfg3"; $B --config cfg3 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json
echo "== cfg1"; $B --config cfg1 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json
