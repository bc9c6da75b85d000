timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
J='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d["n_gpus"], round(d["ms_per_step"],3), round(d["value"],1), "e2e", d["e2e"])'
python bench.py --steps 20 --warmup 5 --no-cpu-baseline | python -c "$J"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 20 --warmup 5 2>gpurun_out/exp21.err | python -c "$J"
tail -3 gpurun_out/exp21.err
