# usage: tools/scale.sh <tag> "<N N ...>" [config] [steps]  (GPU box with >= max N GPUs): the strong-scaling bench line of a config at each N,
# through the library's own exchange (vrt_comm_*). Output: gpurun_out/<tag>_<config>_N<n>.json
T=$1; NS=$2; C=${3:-cfg4}; K=${4:-20}
for n in $NS; do
  if [ $n -eq 1 ]; then timeout 600 python bench.py --gpus 1 --config $C --steps $K --warmup 5 --no-cpu-baseline > gpurun_out/${T}_${C}_N$n.json 2> gpurun_out/${T}_${C}_N$n.err
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --config $C --steps $K --warmup 5 > gpurun_out/${T}_${C}_N$n.json 2> gpurun_out/${T}_${C}_N$n.err; fi
  echo "N=$n rc=$?"; python - gpurun_out/${T}_${C}_N$n.json <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(d["n_gpus"], "ms/frame", round(d["ms_per_step"], 3), "Mrays/s", round(d["value"], 1), "e2e ms", round(d["e2e"]["ms_per_step"], 3), "latency", round(d["latency_ms_single_frame"], 3), {k: round(v, 3) for k, v in d["roofline"]["stages_ms"].items()}, "launches", d["gpu_launches"])
except Exception as e:
    print("parse failed", e); import subprocess; print(subprocess.run(["tail", "-5", sys.argv[1].replace(".json", ".err")], capture_output=True, text=True).stdout)
PY
done
