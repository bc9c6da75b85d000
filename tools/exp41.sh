B="python bench.py --no-cpu-baseline --steps 12 --warmup 3"
get() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(round(d["ms_per_step"],3), "e2e", round(d["e2e"]["ms_per_step"],3), d["roofline"].get("stages_ms"))
PY
}
for f in 2 3 4; do echo "== full F=$f"; $B --frames-in-flight $f > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; done
echo "== full serial"; $B --serial-frames > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json
for f in 2 3 4; do echo "== shard 0/8 F=$f"; $B --shard 0 8 --frames-in-flight $f > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; done
for f in 2 3; do echo "== shard 0/8 no side overlap F=$f"; VRT_OVERLAP=0 $B --shard 0 8 --frames-in-flight $f > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; done
echo "== cfg1"; $B --config cfg1 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json
echo "== cfg2"; $B --config cfg2 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json
echo "== cfg3"; $B --config cfg3 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json
tail -3 gpurun_out/x.err
