# usage: tools/shard_exp.sh : the work of ONE rank of an 8-GPU run (tile shard 0 of 8, no exchange) on one GPU under measurement knobs
run() { echo "== $1"; shift; env "$@" timeout 300 python bench.py --shard 0 8 --steps 20 --warmup 5 --min-timed-ms 300 --no-cpu-baseline $ARGS 2>gpurun_out/shard_exp.err | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ms/frame', round(d['ms_per_step'],4), 'latency', round(d['latency_ms_single_frame'],4), {k: round(v,3) for k,v in d['roofline']['stages_ms'].items()})
except Exception as e: print('failed', e)
" || tail -3 gpurun_out/shard_exp.err; }
ARGS=""
run "default" A=1
run "chain from wave 1" VRT_BOUNCE_CHAIN=1
run "chain from wave 0" VRT_BOUNCE_CHAIN=0
ARGS="--frames-in-flight 4"
run "4 frames in flight" A=1
ARGS="--frames-in-flight 5"
run "5 frames in flight" A=1
run "5 frames, chain 1" VRT_BOUNCE_CHAIN=1
