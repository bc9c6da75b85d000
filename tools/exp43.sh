for n in 2; do python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 40 --warmup 5 > gpurun_out/h$n.json 2> gpurun_out/h$n.err; python - gpurun_out/h$n.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(d["n_gpus"], round(d["ms_per_step"],3), "e2e", round(d["e2e"]["ms_per_step"],3), "host submit", round(d["host_submit_ms_per_step"],3))
PY
done
python bench.py --steps 40 --warmup 5 --no-cpu-baseline --shard 0 8 > gpurun_out/h1.json 2> gpurun_out/h1.err; python - gpurun_out/h1.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print("shard 0/8", round(d["ms_per_step"],3), "host submit", round(d["host_submit_ms_per_step"],3))
PY
