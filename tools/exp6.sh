# 2-GPU bench through torchrun (NCCL gather of colour shards on rank 0), and the reference arm
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3
echo "== reference arm"
python bench.py --impl reference --steps 2 --warmup 1
echo "== full default bench N=1"
time python bench.py
