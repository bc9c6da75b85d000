# usage: tools/e2e_exp.sh <N> "<ENV=val ...>" [bench args]: device-timed and end-to-end ms/frame at N ranks under environment knobs
N=$1; E="$2"; shift 2
for e in $E; do
  [ "$e" = "-" ] && e="A=1"
  echo "== $e $@"
  env $e timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 20 --warmup 5 --min-timed-ms 400 "$@" 2>gpurun_out/e2e_exp.err | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ms/frame', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['ms_per_step'],3), 'launches', d['gpu_launches'])
except Exception as e: print('failed', e)
" || tail -3 gpurun_out/e2e_exp.err
done
