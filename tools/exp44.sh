python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for ex in gather push; do python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 40 --warmup 5 --exchange $ex > gpurun_out/h_$ex.json 2> gpurun_out/h_$ex.err; echo "rc=$?"; tail -3 gpurun_out/h_$ex.err | cut -c1-300; python - gpurun_out/h_$ex.json $ex <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[2], d["n_gpus"], round(d["ms_per_step"],3), "e2e", round(d["e2e"]["ms_per_step"],3), "host submit", round(d["host_submit_ms_per_step"],3))
except Exception as e: print("no result", e)
PY
done
