# reflection chain: from which wave the rest of every chain runs in place, on a 1/8 shard (8-GPU regime) and on the full frame
show() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[2].ljust(34), "ms", round(d["ms_per_step"],4), {k: round(v,3) for k,v in d["roofline"]["stages_ms"].items()})
except Exception as e: print(sys.argv[2], "no result", e)
PY
}
for bc in 2 1 0; do VRT_BOUNCE_CHAIN=$bc python bench.py --shard 0 8 --no-cpu-baseline --steps 60 --warmup 5 > gpurun_out/e48_s_$bc.json 2> gpurun_out/e48.err; show gpurun_out/e48_s_$bc.json "shard 1/8 chain=$bc"; done
for f in 2 4 6; do python bench.py --shard 0 8 --no-cpu-baseline --steps 60 --warmup 5 --frames-in-flight $f > gpurun_out/e48_f_$f.json 2>> gpurun_out/e48.err; show gpurun_out/e48_f_$f.json "shard 1/8 frames-in-flight=$f"; done
VRT_BOUNCE_CHAIN=1 python bench.py --no-cpu-baseline --steps 20 --warmup 3 > gpurun_out/e48_full_1.json 2>> gpurun_out/e48.err; show gpurun_out/e48_full_1.json "full chain=1"
tail -2 gpurun_out/e48.err
