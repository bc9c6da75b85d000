B="python bench.py --no-cpu-baseline --steps 8 --warmup 3"
get() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(round(d["ms_per_step"],3), d["roofline"].get("stages_ms"))
PY
}
for rep in 1 2; do for v in base pf; do echo "== $v full"; VRT_LIB=tools/libvrt_$v.so $B > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; done; done
for v in base pf; do echo "== $v shard 0/8"; VRT_LIB=tools/libvrt_$v.so $B --shard 0 8 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; done
for v in base pf; do echo "== $v cfg3"; VRT_LIB=tools/libvrt_$v.so $B --config cfg3 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; done
VRT_LIB=tools/libvrt_pf.so python -m pytest tests -m gpu -x -q 2>&1 | tail -2
