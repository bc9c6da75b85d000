# 2 GPUs: the frame exchange as a gather + de-tile (default) and as direct tile stores into rank 0's frame over NVLink (--exchange push)
for ex in gather push; do python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 40 --warmup 5 --exchange $ex > gpurun_out/r1_exchange_2gpu_$ex.json 2> gpurun_out/r1_exchange_$ex.err; echo "rc=$?"; tail -2 gpurun_out/r1_exchange_$ex.err | cut -c1-300; python - gpurun_out/r1_exchange_2gpu_$ex.json $ex <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[2], d["n_gpus"], "ms", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["ms_per_step"],3), "host submit", round(d.get("host_submit_ms_per_step",0),3), d["config"].get("exchange"), d["config"].get("parallelism"))
except Exception as e: print("no result", e)
PY
done
