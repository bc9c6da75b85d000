N=$1; shift
run() { echo "== $1"; shift; env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 20 --warmup 5 --min-timed-ms 400 $ARGS 2>gpurun_out/e2e_exp.err | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ms/frame', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['ms_per_step'],3), 'launches', d['gpu_launches'])
except Exception as e: print('failed', e)
" || tail -3 gpurun_out/e2e_exp.err; }
ARGS="$@"
run "host frame by push kernel" A=1
run "host frame fused into the pass" VRT_COMM_FUSE_HOST=1
