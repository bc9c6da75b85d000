B="python bench.py --steps 10 --warmup 3 --no-cpu-baseline"
J='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d["roofline"]; print(round(d["ms_per_step"],3), round(d["value"],1), r.get("stages_ms"))'
echo "== default"; $B | python -c "$J"
echo "== shard 0/8"; $B --shard 0 8 | python -c "$J"
echo "== shard 3/8"; $B --shard 3 8 | python -c "$J"
echo "== shard 0/2"; $B --shard 0 2 | python -c "$J"
