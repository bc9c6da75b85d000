timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
B="python bench.py --steps 10 --warmup 3 --no-cpu-baseline"
J='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d["roofline"]; print(round(d["ms_per_step"],3), round(d["value"],1), r.get("stages_ms"))'
echo "== default (any-hit early-out, levels to 512)"; $B | python -c "$J"
echo "== engine off"; VRT_ENGINE_T=0 $B | python -c "$J"
echo "== traversal 4"; $B --traversal 4 | python -c "$J"
echo "== ao only"; $B --passes 21 | python -c "$J"
echo "== shadow only"; $B --passes 19 | python -c "$J"
echo "== shard 0/8"; $B --shard 0 8 | python -c "$J"
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/shard8_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --shard 0 8 > /dev/null 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open("gpurun_out/shard8_launches.csv")) if len(r)>10]
h=rows[0]; ik=h.index("Kernel Name"); iv=h.index("Metric Value")
last=rows[-20:]
for r in last: print(r[ik][:60], r[iv])
PY
