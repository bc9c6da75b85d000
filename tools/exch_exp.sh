# usage: tools/exch_exp.sh <N> [bench args]: the exchange variants against the exchange-free ideal at N ranks (measurement knobs are env vars)
N=$1; shift
run() { echo "== $1"; shift; env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 20 --warmup 5 --min-timed-ms 400 $ARGS 2>gpurun_out/exch_exp.err | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ms/frame', round(d['ms_per_step'],3), 'region', [round(x,2) for x in d['run']['region_ms_min_median_max']], 'e2e', round(d['e2e']['ms_per_step'],3), 'host', round(d['host_submit_ms_per_step'],3), 'launches', d['gpu_launches'])
except Exception as e: print('failed', e)
" || tail -3 gpurun_out/exch_exp.err; }
ARGS="$@"
run "default (memops + push stream)" A=1
run "no exchange (ideal)" VRT_BENCH_NOEXCHANGE=1
run "round-2a behaviour (kernels on the frame stream)" VRT_COMM_NO_MEMOPS=1 VRT_COMM_PUSH_ON_FRAME_STREAM=1
run "memops, push on frame stream" VRT_COMM_PUSH_ON_FRAME_STREAM=1
