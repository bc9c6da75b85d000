# usage: tools/quick.sh <tag> [bench args...]  (on the GPU box via gpurun): the fast parity subset, then one bench line without the CPU legs,
# then the engine statistics of the same frames (stats build, if present). Output under gpurun_out/<tag>_*.
T=$1; shift
timeout 600 python -m pytest tests/test_gpu.py -q -x -k "cfg4_cfg5 or production or degenerate or ragged or cfg3" 2>&1 | tail -3
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err || tail -5 gpurun_out/${T}_bench.err
python - gpurun_out/${T}_bench.json <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r = d["roofline"]
print("ms/frame", round(d["ms_per_step"], 3), "latency", round(d["latency_ms_single_frame"], 3), "e2e", round(d["e2e"]["ms_per_step"], 3), "stages", {k: round(v, 3) for k, v in r["stages_ms"].items()}, "frac", round(r["frac"], 3), "issue", round(r["issue"]["frac"], 3))
PY
if [ -f voxel-raytracing_b200/libvrt_stats.so ]; then
  VRT_LIB=$PWD/voxel-raytracing_b200/libvrt_stats.so VRT_DEBUG_STATS=1 timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline "$@" 2>&1 >/dev/null | grep "engine stats" | tail -1
fi
