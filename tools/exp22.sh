B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline"
J='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d["roofline"]; print(round(d["ms_per_step"],3), round(d["value"],1), r.get("stages_ms"))'
for i in 1 2; do
echo "== shard 0/8 overlap"; $B --shard 0 8 | python -c "$J"
echo "== shard 0/8 no overlap"; VRT_NO_OVERLAP=1 $B --shard 0 8 | python -c "$J"
echo "== full overlap"; $B | python -c "$J"
echo "== full no overlap"; VRT_NO_OVERLAP=1 $B | python -c "$J"
done
