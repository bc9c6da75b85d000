B="python bench.py --no-cpu-baseline --steps 8 --warmup 3"
get() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(round(d["ms_per_step"],3), d["roofline"].get("stages_ms"))
PY
}
for v in "" tools/libvrt_u8.so tools/libvrt_u16.so; do echo "== lib $v"; if [ -n "$v" ]; then export VRT_LIB=$v; fi; $B > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; $B --passes 1 > gpurun_out/x.json 2>gpurun_out/x.err; get gpurun_out/x.json; done
unset VRT_LIB
bash tools/profile_round.sh r1
