B="python bench.py --no-cpu-baseline --steps 8 --warmup 3"
get() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(round(d["ms_per_step"],3), d["roofline"].get("stages_ms"))
PY
}
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for v in n4f16 n8f16 n2f16 n4f8; do for fs in 2 4 6; do echo "== $v farshare $fs"; VRT_ENGINE_FARSHARE=$fs VRT_LIB=tools/libvrt_$v.so $B > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; done; done
for v in n4f16 n8f16; do echo "== cfg3 $v"; VRT_LIB=tools/libvrt_$v.so $B --config cfg3 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; done
