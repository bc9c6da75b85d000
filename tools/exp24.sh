timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline"
J='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d["roofline"]; print(round(d["ms_per_step"],3), round(d["value"],1), r.get("stages_ms"))'
echo "== default"; $B | python -c "$J"
echo "== shard 0/8"; $B --shard 0 8 | python -c "$J"
echo "== T 12"; VRT_ENGINE_T=12 $B | python -c "$J"
echo "== T 6"; VRT_ENGINE_T=6 $B | python -c "$J"
