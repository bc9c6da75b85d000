timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
B="python bench.py --steps 10 --warmup 3 --no-cpu-baseline"
J='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d["roofline"]; print(round(d["ms_per_step"],3), round(d["value"],1), r.get("stages_ms"))'
echo "== default (overlap)"; $B | python -c "$J"
echo "== no overlap"; VRT_NO_OVERLAP=1 $B | python -c "$J"
echo "== shard 0/8"; $B --shard 0 8 | python -c "$J"
echo "== shard 0/8 no overlap"; VRT_NO_OVERLAP=1 $B --shard 0 8 | python -c "$J"

