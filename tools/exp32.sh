python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/p1.json 2> gpurun_out/p1.err; echo rc=$?; tail -2 gpurun_out/p1.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/p2.json 2> gpurun_out/p2.err; echo rc=$?; tail -3 gpurun_out/p2.err
for f in gpurun_out/p1.json gpurun_out/p2.json; do python - $f <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(d["n_gpus"], round(d["ms_per_step"],3), round(d["value"],1), "e2e ms", d["e2e"]["ms_per_step"], d["gpu_launches"], d["roofline"].get("stages_ms"))
PY
done
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
