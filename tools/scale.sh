# strong scaling of the cfg4 frame over N GPUs of one box (tiles interleaved over ranks, colour shards gathered on rank 0 over NCCL)
for n in 1 2 4 8; do
  if [ $n -eq 1 ]; then python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err; fi
  echo "N=$n rc=$?"; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/scale_$n.json").read().strip().splitlines()[-1]); print(d["n_gpus"], round(d["ms_per_step"],3), round(d["value"],1), d["e2e"]["ms_per_step"], d["roofline"].get("stages_ms"))
except Exception as e: print("parse failed", e)
PY
done
# cfg5: BASELINE's 8-GPU line — 2048^3 volume replicated per GPU, 7680x4320, 256-frame camera path, colour shards gathered on rank 0
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 bench.py --config cfg5 --gpus 8 --steps 256 --warmup 3 > gpurun_out/scale_cfg5_8.json 2> gpurun_out/scale_cfg5_8.err
echo "cfg5 N=8 rc=$?"; cut -c1-400 gpurun_out/scale_cfg5_8.json; tail -2 gpurun_out/scale_cfg5_8.err
