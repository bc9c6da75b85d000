B="python bench.py --no-cpu-baseline --steps 6 --warmup 3"
for r in 0 1 2 3 4 5 6 7; do $B --shard $r 8 > gpurun_out/x.json 2>gpurun_out/x.err; python - gpurun_out/x.json $r <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print("shard", sys.argv[2], round(d["ms_per_step"],3), round(d["rays_per_frame"]/1e6,3), {k: round(v,3) for k,v in d["roofline"]["stages_ms"].items()})
PY
done
